"""Host-side mirror of the reference's ``models/nerf.py`` (PosEmbedding :6-32, NeRF :80-212).

Same constructor arguments, attribute names and ``state_dict`` keys, so reference
checkpoints load unchanged and ``train.py`` / ``eval.py`` can import these classes instead.  The
modules hold parameters only: every forward runs inside libnerfw_b200.so (tcgen05 field kernels).
There is no PyTorch fallback -- CPU tensors raise.
"""
from __future__ import annotations

import torch
from torch import nn

from . import _lib as L
from . import kernels as K

__all__ = ["PosEmbedding", "NeRF"]


class PosEmbedding(nn.Module):
    """(B,3) -> (B, 6*N_freqs+3): [x, sin(2^k x), cos(2^k x), ...]  (models/nerf.py:6-32).

    Only the reference's default ``logscale=True`` with ``max_logscale == N_freqs-1`` (octave
    frequencies 2^k, the only setting train.py:46-47 / eval.py:121-122 ever build) is supported:
    the kernels rely on the frequencies being exact powers of two."""

    def __init__(self, max_logscale, N_freqs, logscale=True):
        super().__init__()
        if not logscale or max_logscale != N_freqs - 1:
            raise NotImplementedError(
                "nerf_fl_b200.PosEmbedding supports octave frequencies only "
                "(logscale=True, max_logscale == N_freqs-1)")
        self.N_freqs = int(N_freqs)
        self.freqs = 2 ** torch.linspace(0, max_logscale, N_freqs)   # plain attribute, as in nerf.py:15

    def forward(self, x):
        return K.posenc(x, self.N_freqs)


class NeRF(nn.Module):
    """Parameter container with the reference's layout (models/nerf.py:81-151) + fused forward."""

    def __init__(self, typ, D=8, W=256, skips=[4], in_channels_xyz=63, in_channels_dir=27,
                 encode_appearance=False, in_channels_a=48, encode_transient=False, in_channels_t=16,
                 beta_min=0.03, refine_pose=False):
        super().__init__()
        if D != 8 or W != 256 or list(skips) != [4]:
            raise NotImplementedError("nerf_fl_b200.NeRF is specialised for D=8, W=256, skips=[4] "
                                      "(the only shape the reference builds, train.py:60-70)")
        if refine_pose:
            raise NotImplementedError("refine_pose (BARF) is outside the hot path (SURVEY.md section 2, row 2b)")
        self.typ = typ
        self.D, self.W, self.skips = D, W, skips
        self.in_channels_xyz = in_channels_xyz
        self.in_channels_dir = in_channels_dir
        self.refine_pose = False
        self.encode_appearance = False if typ == "coarse" else encode_appearance
        self.in_channels_a = in_channels_a if encode_appearance else 0
        self.encode_transient = False if typ == "coarse" else encode_transient
        self.in_channels_t = in_channels_t
        self.beta_min = beta_min

        for i in range(D):
            fan_in = in_channels_xyz if i == 0 else (W + in_channels_xyz if i in skips else W)
            setattr(self, f"xyz_encoding_{i + 1}", nn.Sequential(nn.Linear(fan_in, W), nn.ReLU(True)))
        self.xyz_encoding_final = nn.Linear(W, W)
        self.dir_encoding = nn.Sequential(nn.Linear(W + in_channels_dir + self.in_channels_a, W // 2), nn.ReLU(True))
        self.static_sigma = nn.Sequential(nn.Linear(W, 1), nn.Softplus())
        self.static_rgb = nn.Sequential(nn.Linear(W // 2, 3), nn.Sigmoid())
        if self.encode_transient:
            self.transient_encoding = nn.Sequential(
                nn.Linear(W + in_channels_t, W // 2), nn.ReLU(True), nn.Linear(W // 2, W // 2), nn.ReLU(True),
                nn.Linear(W // 2, W // 2), nn.ReLU(True), nn.Linear(W // 2, W // 2), nn.ReLU(True))
            self.transient_sigma = nn.Sequential(nn.Linear(W // 2, 1), nn.Softplus())
            self.transient_rgb = nn.Sequential(nn.Linear(W // 2, 3), nn.Sigmoid())
            self.transient_beta = nn.Sequential(nn.Linear(W // 2, 1), nn.Softplus())
        self._nwb_cache = None      # (key, packed operand cache)
        self._nwb_dirty = False     # set by FlatTrainState.adam_step (updates bypass torch's version counters)

    # ---- kernel-side views ------------------------------------------------------------------
    def nwb_desc(self) -> L.FieldDesc:
        # the a-columns of dir_encoding exist iff the ctor was given encode_appearance (nerf.py:109,
        # also for typ == 'coarse', where the renderer never feeds them)
        return K.field_desc(self.in_channels_xyz, self.in_channels_dir, self.in_channels_a,
                            self.in_channels_t if self.encode_transient else 0)

    def nwb_params(self) -> dict:
        """name -> parameter tensor (fp32, CUDA, contiguous), reference key names."""
        sd = dict(self.named_parameters())
        for k, v in sd.items():
            if not (v.is_cuda and v.dtype == torch.float32 and v.is_contiguous()):
                raise RuntimeError(f"nerf_fl_b200.NeRF parameter {k} must be a contiguous fp32 CUDA tensor "
                                   "(there is no CPU path)")
        return sd

    def nwb_packed(self) -> torch.Tensor:
        """bf16 tile-image weight cache; rebuilt (one kernel) whenever a parameter changed."""
        sd = self.nwb_params()
        key = tuple((v.data_ptr(), v._version) for v in sd.values())
        if self._nwb_cache is None or self._nwb_cache[0] != key or self._nwb_dirty:
            dev = next(iter(sd.values())).device
            out = self._nwb_cache[1] if self._nwb_cache is not None and self._nwb_cache[1].device == dev else None
            packed = K.field_pack(self.nwb_desc(), {k: v.detach() for k, v in sd.items()}, out)
            self._nwb_cache = (key, packed)
            self._nwb_dirty = False
        return self._nwb_cache[1]

    def forward(self, x, sigma_only=False, output_transient=True):
        """models/nerf.py:153-212 on already-embedded rows x (B, 63 | 90(+48) | +16)."""
        from .rendering import field_apply
        return field_apply(self, x, sigma_only, output_transient)

    def extra_repr(self):
        return f"typ={self.typ}, encode_appearance={self.encode_appearance}, encode_transient={self.encode_transient}"


def field_macs(model: NeRF, sigma_only: bool = False, output_transient: bool | None = None) -> int:
    """Multiply-accumulates per sample of one forward, used by bench.py."""
    W, ix, idr = model.W, model.in_channels_xyz, model.in_channels_dir
    trunk = ix * W + 3 * W * W + (W + ix) * W + 3 * W * W
    if sigma_only:
        return trunk + W
    ia = model.in_channels_a if model.encode_appearance else 0
    m = trunk + W + W * W + (W + idr + ia) * (W // 2) + 3 * (W // 2)
    tr = model.encode_transient if output_transient is None else (output_transient and model.encode_transient)
    if tr:
        m += (W + model.in_channels_t) * (W // 2) + 3 * (W // 2) ** 2 + 5 * (W // 2)
    return m
