"""Drop-in for the reference's model/mdm_old.py `MDM_Old` ("MDM V1": plain MDM trans_enc).

Constructor signature, attribute names and state_dict keys follow model/mdm_old.py:7-67 (same
sub-module creation order -> same random init under the same seed); `forward` (:84-122) runs on the
B200 engine.  This is the only reference path that accepts T = 196 (HumanML3D / KIT shapes).
"""
from .base import (EngineBackedDenoiser, InputProcess, OutputProcess, PositionalEncoding, Rotation2xyz,
                   SeedPoseEncoder, TimestepEmbedder, make_encoder)


class MDM_Old(EngineBackedDenoiser):
    arch = "v1"

    def __init__(self, njoints, nfeats, translation, pose_rep, glob, glob_rot, latent_dim=256, ff_size=1024,
                 num_layers=8, num_heads=4, dropout=0.1, activation="gelu", data_rep='rot6d', dataset='amass',
                 **kargs):
        super().__init__()
        self.dataset, self.pose_rep = dataset, pose_rep
        self.njoints, self.nfeats = njoints, nfeats
        self.input_feats = njoints * nfeats
        self.latent_dim = latent_dim
        self.cond_mode = kargs.get('cond_mode', 'no_cond')
        self.glob, self.glob_rot, self.translation = glob, glob_rot, translation
        if data_rep != 'genea_vec':
            raise NotImplementedError("only data_rep='genea_vec' exists on the reference path")
        if activation != "gelu":
            raise NotImplementedError("the B200 FFN epilogue implements erf-GELU")

        self.cond_mask_prob = kargs.get('cond_mask_prob', 0.)
        self.seed_poses = kargs.get('seed_poses', 0)
        assert self.seed_poses > 0
        self.seed_pose_encoder = SeedPoseEncoder(njoints, self.seed_poses, latent_dim)

        self.mfcc_dim = 26
        self.data_rep = data_rep
        self.input_process = InputProcess(data_rep, self.input_feats + self.mfcc_dim, latent_dim)

        self.num_heads, self.ff_size, self.dropout = num_heads, ff_size, dropout
        self.activation, self.num_layers = activation, num_layers
        self.seqTransEncoder = make_encoder(latent_dim, num_heads, ff_size, dropout, activation, num_layers)

        self.sequence_pos_encoder = PositionalEncoding(latent_dim, dropout)
        self.embed_timestep = TimestepEmbedder(latent_dim, self.sequence_pos_encoder)
        self.output_process = OutputProcess(data_rep, self.input_feats, latent_dim, njoints, nfeats)
        self.rot2xyz = Rotation2xyz(device='cpu', dataset=dataset)
