"""Policy/value network: PyTorch definition (weights, training, CPU reference) and the
B200 inference engine behind it.

`Net` keeps the reference's constructor and state_dict layout (ai/nn.py:63-72) so that
checkpoints written by utils.save_model load unchanged; `img_size` may also be a
(rows, cols) pair -- the reference only supports square boards (ai/nn.py:26), Connect4's
6x7 is a build extension.  `B200Net` is the inference path used by self-play: it uploads a
state_dict to libaz_b200.so (tcgen05 GEMMs + fused kernels, csrc/az_net.cu).
"""
import ctypes as C

import numpy as np
import torch
from torch import nn

from .. import lib as _lib


def _pair(img_size):
    if isinstance(img_size, (tuple, list)):
        return int(img_size[0]), int(img_size[1])
    return int(img_size), int(img_size)


class _Conv(nn.Module):                      # state_dict: embedding.patch_embed.patch_embed.*
    def __init__(self, channels, dim, k):
        super().__init__()
        self.patch_embed = nn.Conv2d(channels, dim, kernel_size=k, stride=1, padding=k // 2)

    def forward(self, boards):
        return self.patch_embed(boards).flatten(2).transpose(1, 2)       # [B, H*W, dim]


class _Tokens(nn.Module):                    # state_dict: embedding.{cls_token,pos_embedding}
    def __init__(self, rows, cols, k, dim, channels):
        super().__init__()
        self.patch_embed = _Conv(channels, dim, k)
        self.cls_token = nn.Parameter(torch.randn(1, 1, dim))
        self.pos_embedding = nn.Parameter(torch.randn(1, rows * cols + 1, dim))

    def forward(self, boards):
        tok = self.patch_embed(boards)
        cls = self.cls_token.expand(tok.shape[0], -1, -1)
        return torch.cat((cls, tok), dim=1) + self.pos_embedding


class _Block(nn.Module):                     # pre-LN transformer block, ai/nn.py:38-61
    def __init__(self, dim, heads, dropout):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = nn.MultiheadAttention(dim, heads, dropout=dropout, batch_first=True)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = nn.Sequential(nn.Linear(dim, 4 * dim), nn.GELU(), nn.Dropout(dropout),
                                 nn.Linear(4 * dim, dim), nn.Dropout(dropout))

    def forward(self, x):
        h = self.norm1(x)
        x = x + self.attn(h, h, h, need_weights=False)[0]
        return x + self.mlp(self.norm2(x))


class Net(nn.Module):
    """Net(img_size, patch_size, embed_dim, action_dim, num_heads, depth, channels, dropout)."""

    def __init__(self, img_size, patch_size, embed_dim, action_dim, num_heads, depth, channels, dropout=0.0):
        super().__init__()
        rows, cols = _pair(img_size)
        self.config = dict(rows=rows, cols=cols, patch=patch_size, dim=embed_dim, actions=action_dim,
                           heads=num_heads, depth=depth, channels=channels)
        self.embedding = _Tokens(rows, cols, patch_size, embed_dim, channels)
        self.blocks = nn.ModuleList(_Block(embed_dim, num_heads, dropout) for _ in range(depth))
        self.norm = nn.LayerNorm(embed_dim)
        self.policy_head = nn.Linear(embed_dim, action_dim)
        self.value_head = nn.Linear(embed_dim, 1)

    def forward(self, boards):
        x = self.embedding(boards)
        for blk in self.blocks:
            x = blk(x)
        cls = self.norm(x)[:, 0]
        return self.policy_head(cls), torch.tanh(self.value_head(cls))


class B200Net:
    """Inference engine for a `Net` on the GPU.  Callable like the model in ai/mcts.py:46:
    `logits, values = engine(boards)` with boards float32 [B, C, H, W] (CPU tensor/array)."""

    def __init__(self, model_or_config, state_dict=None, max_batch=4096):
        if isinstance(model_or_config, nn.Module):
            cfg = dict(model_or_config.config)
            state_dict = model_or_config.state_dict()
        else:
            cfg = dict(model_or_config)
        self.cfg = cfg
        self._L = _lib.load()
        h = C.c_void_p()
        _lib.check(self._L.az_net_create(cfg["rows"], cfg["cols"], cfg["channels"], cfg["patch"], cfg["dim"],
                                         cfg["heads"], cfg["depth"], cfg["actions"], int(max_batch), C.byref(h)),
                   "az_net_create")
        self.h = h
        if state_dict is not None:
            self.load_state_dict(state_dict)

    def load_state_dict(self, state_dict):
        """Upload weights (fp32 state_dict of `Net`, e.g. after a training update)."""
        for name, t in state_dict.items():
            a = np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy())
            _lib.check(self._L.az_net_set_param(self.h, name.encode(), a.ctypes.data_as(C.c_void_p), a.size),
                       f"az_net_set_param({name})")
        _lib.check(self._L.az_net_finalize(self.h), "az_net_finalize")

    def forward_numpy(self, boards):
        b = np.ascontiguousarray(boards, dtype=np.float32)
        B, A = b.shape[0], self.cfg["actions"]
        logits = np.empty((B, A), np.float32)
        values = np.empty((B, 1), np.float32)
        priors = np.empty((B, A), np.float32)
        _lib.check(self._L.az_net_forward_planes(self.h, b.ctypes.data_as(C.c_void_p), B,
                                                 logits.ctypes.data_as(C.c_void_p), values.ctypes.data_as(C.c_void_p),
                                                 priors.ctypes.data_as(C.c_void_p), None), "az_net_forward_planes")
        return logits, values, priors

    def __call__(self, boards):
        if isinstance(boards, torch.Tensor):
            boards = boards.detach().cpu().numpy()
        logits, values, _ = self.forward_numpy(boards)
        return torch.from_numpy(logits), torch.from_numpy(values)

    def eval_leaves(self, forest, stream=0):
        _lib.check(self._L.az_net_eval_leaves(self.h, forest.h, stream), "az_net_eval_leaves")

    def profile(self, enable=True):
        _lib.check(self._L.az_net_profile(self.h, int(enable)), "az_net_profile")

    def profile_on(self, on):
        """Pause / resume the timing without clearing the totals."""
        _lib.check(self._L.az_net_profile(self.h, 2 if on else 3), "az_net_profile")

    def profile_read(self):
        """{kernel kind: (total ms, launches)} of the eager passes since profile(True)."""
        names = (C.c_char_p * 16)()
        ms = (C.c_double * 16)()
        cnt = (C.c_longlong * 16)()
        k = _lib.check(self._L.az_net_profile_read(self.h, 16, names, ms, cnt), "az_net_profile_read")
        return {names[i].decode(): (ms[i], cnt[i]) for i in range(k)}

    def proj_width(self):
        return int(self._L.az_net_proj_width(self.h))

    def path_flags(self):
        """bit 0 fused front end, 1 reassociated front end, 2 CTA-pair cls-row GEMMs, 3 LayerNorm fold, 4 head epilogue."""
        return int(self._L.az_net_path_flags(self.h))

    def launch_count(self):
        return int(self._L.az_net_launch_count(self.h))

    def close(self):
        if getattr(self, "h", None):
            self._L.az_net_destroy(self.h)
            self.h = None

    __del__ = close
