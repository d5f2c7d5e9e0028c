"""Drop-in for the reference's model/mdm.py `MDM` ("MDM V2": local attention + RoPE + trans_enc).

Constructor signature, attribute names and state_dict keys follow model/mdm.py:11-103 (sub-modules are
created in the same order, so the same torch seed gives the same random init); `forward` (:105-224)
runs on the B200 engine.  Out of scope and rejected loudly: WavEncoder (dead in the reference,
mdm.py:137) and CLIP text encoding -- with `use_text` the caller passes the 512-d embedding tensor as
y['text'] (BASELINE.json config 2: synthetic CLIP-like cond embedding).
"""
import torch.nn as nn

from .base import (EngineBackedDenoiser, InputProcess, OutputProcess, PositionalEncoding, Rotation2xyz,
                   SeedPoseEncoder, SinusoidalEmbeddings, TimestepEmbedder, make_encoder)


class MDM(EngineBackedDenoiser):
    arch = "v2"

    def __init__(self, njoints, nfeats, pose_rep, data_rep, latent_dim=256, text_dim=64, ff_size=1024,
                 num_layers=8, num_heads=4, dropout=0.1, activation="gelu", dataset='amass', clip_dim=512,
                 clip_version=None, **kargs):
        super().__init__()
        self.dataset, self.pose_rep, self.data_rep = dataset, pose_rep, data_rep
        self.njoints, self.nfeats = njoints, nfeats
        self.input_feats = njoints * nfeats
        self.latent_dim, self.dropout = latent_dim, dropout
        if data_rep != 'genea_vec':
            raise NotImplementedError("only data_rep='genea_vec' exists on the reference path (mdm.py:354-358)")

        self.sequence_pos_encoder = PositionalEncoding(latent_dim, dropout)
        self.embed_timestep = TimestepEmbedder(latent_dim, self.sequence_pos_encoder)

        self.use_text = kargs.get('use_text', False)
        self.cond_mask_prob = kargs.get('cond_mask_prob', 0.)
        self.text_dim, self.clip_dim = text_dim, clip_dim
        if self.use_text:
            self.embed_text = nn.Linear(clip_dim, text_dim)
            self.clip_version = clip_version        # CLIP itself is not loaded: y['text'] is the embedding

        self.seed_poses = kargs.get('seed_poses', 0)
        if self.seed_poses <= 0:
            raise NotImplementedError("the reference forward requires seed poses (mdm.py:125-127)")
        self.seed_pose_encoder = SeedPoseEncoder(njoints, self.seed_poses,
                                                 latent_dim - text_dim if self.use_text else latent_dim)

        self.mfcc_input = kargs.get('mfcc_input', False)
        self.use_wav_enc = kargs.get('use_wav_enc', False)
        if not self.mfcc_input or self.use_wav_enc:
            raise NotImplementedError("only mfcc_input=True is runnable in the reference (mdm.py:133-141)")
        self.mfcc_dim = 26
        self.audio_feat_dim = self.mfcc_dim

        self.input_process = InputProcess(data_rep, self.input_feats, latent_dim)

        self.cl_head = 8
        self.project_to_lat = nn.Linear(latent_dim * 2 + self.audio_feat_dim, latent_dim)
        self.rel_pos = SinusoidalEmbeddings(latent_dim // self.cl_head)

        self.num_heads, self.ff_size, self.activation, self.num_layers = num_heads, ff_size, activation, num_layers
        if activation != "gelu":
            raise NotImplementedError("the B200 FFN epilogue implements erf-GELU (model_util.py:31)")
        self.seqTransEncoder = make_encoder(latent_dim, num_heads, ff_size, dropout, activation, num_layers)

        self.output_process = OutputProcess(data_rep, self.input_feats, latent_dim, njoints, nfeats)
        self.rot2xyz = Rotation2xyz(device='cpu', dataset=dataset)

    def encode_text(self, raw_text):
        raise NotImplementedError("CLIP is out of scope: pass the text embedding tensor as y['text']")
