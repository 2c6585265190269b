"""Shared helpers for the parity tests."""
import hashlib
import os

import numpy as np

from oracle import portapi, refapi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def plane_md5(planes):
    """Per-plane MD5 of the visible samples (what the reference's framemd5 hashes)."""
    return [md5(np.ascontiguousarray(p)) for p in planes]


def noise_planes(seed, dims, bd=10):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 1 << bd, (h, w)).astype(np.uint16) for (w, h) in dims]


def oracle_residual(planes, pic, want_residual=False):
    """The checker: compiled reference when present (oracle/_ref), else the C port."""
    if refapi.available():
        return refapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], pic["intra_mode"], want_residual)
    return portapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], want_residual)


def to_device_planes(planes_np, torch, align=64):
    """Upload numpy planes into pitch-padded torch CUDA tensors (views of width w)."""
    out = []
    for p in planes_np:
        h, w = p.shape
        stride = (w + align - 1) // align * align
        t = torch.zeros((h, stride), dtype=torch.int16, device="cuda")
        t[:, :w] = torch.from_numpy(p.view(np.int16)).cuda()
        out.append(t[:, :w])
    return out


def from_device_planes(tensors):
    return [t.cpu().numpy().view(np.uint16) for t in tensors]
