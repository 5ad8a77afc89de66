"""CPU restatement of the rep producers in front of the kNN reward (TEST INFRASTRUCTURE ONLY).

  icm_trunk_rep  <- ICM.trunk = nn.Sequential(nn.Linear, nn.LayerNorm, nn.Tanh) and ICM.get_rep,
                    /root/reference/agents/unsupervised_learning/icm_apt.py:17-18,50-52
  l2_normalize   <- F.normalize(z, dim=1, p=2), agents/unsupervised_learning/proto.py:104

Pinned bitwise to tests/golden/rep_reference.npz (outputs of the reference's own ICM class,
oracle/gen_golden.py:gen_rep) by tests/test_oracle_pins.py.
"""
import torch
import torch.nn.functional as F


def icm_trunk_rep(obs, weight, bias, ln_weight, ln_bias, eps=1e-5):
    h = F.linear(obs, weight, bias)                                   # icm_apt.py:17 nn.Linear
    h = F.layer_norm(h, (weight.shape[0],), ln_weight, ln_bias, eps)  # :18 nn.LayerNorm
    return torch.tanh(h)                                              # :18 nn.Tanh


def l2_normalize(x, eps=1e-12):
    # F.normalize(x, dim=1, p=2): x / max(||x||_2, eps)
    return x / x.norm(2, 1, keepdim=True).clamp_min(eps).expand_as(x)
