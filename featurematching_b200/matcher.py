"""The whole matcher on the GPU: the wiring of the reference's `net` (network/net.py:20-102) over the
B200 modules, for boxes where the reference tree is absent.

    Matcher(config)            same lower-case config dict as net(config) (demo/net_config.py:48)
    Matcher.forward(data)      same contract as net.forward (network/net.py:40-92): reads
                               data['image0'], data['image1'] [B, C, H, W]; writes hw*_i/c/f, the coarse
                               and fine match keys, feat_c0/1 and feat_f0/1
    Matcher.load_state_dict    strips the 'matcher.' / 'loftr_' prefixes like network/net.py:94-102

Attribute names (backbone, coarse, coarse_matching, fine_preprocess, fine, fine_matching) and every
parameter name equal the reference's, so `net.state_dict()` loads here and vice versa.  The backbone
is stock PyTorch (backbone.py, library-level); everything from the coarse transformer on runs in
libfm_b200.so.  `install(net)` (modules.py) is the other way in: it swaps the B200 modules into a
constructed reference `net`.
"""
import torch
import torch.nn as nn

from . import modules
from .backbone import SwinUNetBackbone


def default_config():
    """The matcher hyper-parameters the reference ships (demo/net_config.py:8-46) as the lower-cased dict its modules
    receive: the configuration `Matcher`, the bench and the demos run when no other is given."""
    return {
        "fine_window_size": 7,
        "fine_concat_coarse_feat": True,
        "coarse": {"d_model": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4, "attention": "linear"},
        "match_coarse": {"thr": 0.2, "border_rm": 2, "match_type": "dual_softmax", "dsmax_temperature": 0.1,
                         "train_coarse_percent": 1.0, "train_pad_num_gt_min": 200},
        "fine": {"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"], "attention": "linear"},
        "pose_net_flag": "none",
        "backbone_type": "swin", "input_channel": 3, "resolution": (8, 2),
    }


class Matcher(nn.Module):
    def __init__(self, config, emit_bf16_fine=True):
        super().__init__()
        self.config = config
        if config.get('backbone_type', 'swin') != 'swin':
            raise ValueError(f"net.BACKBONE_TYPE {config.get('backbone_type')} not supported.")
        if config.get('pose_net_flag', 'none') in ('old', 'new'):
            raise NotImplementedError("the learnt pose heads (network/utils/pose_pred*.py) are not part of "
                                      "the B200 matcher; pose from matches is featurematching_b200.pose")
        self.backbone = SwinUNetBackbone(in_chans=config.get('input_channel', 3))
        self.coarse = modules.LocalFeatureTransformer(config['coarse'])
        self.coarse_matching = modules.CoarseMatching(config['match_coarse'])
        self.fine_preprocess = modules.FinePreprocess(config)
        self.fine = modules.LocalFeatureTransformer(config['fine'])
        self.fine_matching = modules.FineMatching(config['fine'])
        self.emit_bf16_fine = emit_bf16_fine
        self.train(False)

    def forward(self, data):
        if self.training:
            raise NotImplementedError("Matcher (B200) is inference-only")
        with torch.no_grad():
            data.update({'bs': data['image0'].size(0),
                         'hw0_i': data['image0'].shape[2:], 'hw1_i': data['image1'].shape[2:]})
            feats_c, feats_f = self.backbone(torch.cat([data['image0'], data['image1']], dim=0),
                                             emit_bf16=self.emit_bf16_fine)
            (feat_c0, feat_c1), (feat_f0, feat_f1) = feats_c.split(data['bs']), feats_f.split(data['bs'])
            data.update({'hw0_c': feat_c0.shape[2:], 'hw1_c': feat_c1.shape[2:],
                         'hw0_f': feat_f0.shape[2:], 'hw1_f': feat_f1.shape[2:]})
            # 'n c h w -> n (h w) c' (net.py:69-70): a view, the maps are stored channels-last
            feat_c0 = feat_c0.permute(0, 2, 3, 1).reshape(feat_c0.shape[0], -1, feat_c0.shape[1])
            feat_c1 = feat_c1.permute(0, 2, 3, 1).reshape(feat_c1.shape[0], -1, feat_c1.shape[1])
            feat_c0, feat_c1 = self.coarse(feat_c0, feat_c1, None, None)
            self.coarse_matching(feat_c0, feat_c1, data, mask_c0=None, mask_c1=None)
            w0, w1 = self.fine_preprocess(feat_f0, feat_f1, feat_c0, feat_c1, data)
            if w0.size(0) != 0:
                w0, w1 = self.fine(w0, w1)
            self.fine_matching(w0, w1, data)
            data['feat_c0'], data['feat_c1'] = feat_c0, feat_c1
            data['feat_f0'], data['feat_f1'] = feat_f0, feat_f1

    def load_state_dict(self, state_dict, *args, **kwargs):
        sd = {}
        for k, v in state_dict.items():
            if k.startswith('matcher.'):
                k = k.replace('matcher.', '', 1)
            if k.startswith('loftr_'):
                k = k.replace('loftr_', '', 1)
            sd[k] = v
        return super().load_state_dict(sd, *args, **kwargs)
