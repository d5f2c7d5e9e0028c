"""CPU oracle for the GestureDiffusion / MDM denoising sampling loop.

TEST INFRASTRUCTURE ONLY.  Nothing under ``gesturediffusion_b200/`` may import
this package.  The only permitted importers are ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` -- and there only as the checker / reported CPU baseline,
never as the thing that is shipped or measured as the product.

Parity status: PINNED.  Every function here is a plain restatement (numpy
float64 for the schedule tables, torch CPU ops for the denoiser) of a reference
function, cited file:line in its docstring, and is checked against outputs of
the *unmodified* reference (imported from /root/reference through
``oracle/ref_shims.py`` in the build container) by
``tests/golden/make_golden.py`` -> ``tests/golden/*.npz`` ->
``tests/test_oracle_golden.py``.  The reference itself ships no tests and no
golden vectors, so the fixtures generated from it are the pin.
"""
