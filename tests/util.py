"""Helpers shared by the GPU parity tests (oracle = checker only)."""
import numpy as np
import torch

from oracle import c_oracle as co
from oracle import quant_oracle as qo


def dev(a: np.ndarray) -> torch.Tensor:
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t: torch.Tensor) -> np.ndarray:
    a = t.detach().cpu().numpy()
    if a.dtype == np.int32:
        a = a.view(np.uint32)
    return a


def beq(a: np.ndarray, b: np.ndarray) -> bool:
    return a.shape == b.shape and a.dtype == b.dtype and np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


def rand_weight(fmt: qo.Fmt, K: int, N: int, seed: int):
    """Random packed words (all codes, dead bits set for int3) and scales ~U(0, 0.01)."""
    rng = np.random.default_rng(seed)
    packed = rng.integers(0, 2**32, size=(fmt.n_words(K), N), dtype=np.uint64).astype(np.uint32)
    scale = (rng.random((fmt.n_groups(K), N), dtype=np.float32) * 0.01).astype(fmt.np_dtype)
    return packed, scale


def check_gemv_tolerance(y, packed, scale, x, fmt, out_f32, K=None, what="", group_scaled=False):
    """|y - y64| <= rel*|y64| + 2^-14 * sum|x w| + 2^-24 against the f64 evaluation
    of the SAME dequantized weights."""
    y64, ab = co.gemv_f64(packed, scale, x, fmt, K)
    tol = qo.tolerance(y64, ab, out_f32, group_scaled)
    err = np.abs(y.astype(np.float64).reshape(-1) - y64)
    worst = float((err / tol).max())
    assert worst <= 1.0, f"{what}: error/tolerance = {worst:.3f} (max abs err {err.max():.3e})"
    return err.max(), worst
