"""Quantization scheme description: one dataclass that maps 1:1 onto the reference's four flags
(--dtype, --quantization-mode, --quantization-sym, --quantization-storage-nbit; build.py:19-25,
utils.py:31-36) and onto the BASELINE names q4f16_0 / q3f16_0 / q4f32_0 / q8f16_0."""
from __future__ import annotations

import dataclasses

import torch

# ids shared with include/b200q.h
FMT_IDS = {"q4f16_0": 0, "q3f16_0": 1, "q4f32_0": 2, "q8f16_0": 3, "q4f16_asym": 4, "q4f32_asym": 5,
           "q3f16_u16": 6, "fp4f16_0": 7, "fp4f32_0": 8}
EPILOGUES = {
    "none": 0,
    "add": 1,
    "multiply": 2,
    "silu": 3,
    "cast_f32": 4,
    "bias": 5,
    "bias_gelu": 6,
    "bias_add2": 7,
}
FLAG_PDL = 1


@dataclasses.dataclass(frozen=True)
class QuantSpec:
    name: str
    mode: str  # "int4" | "int3" | "int8"   (--quantization-mode)
    dtype: str  # "float16" | "float32"     (--dtype)
    group_size: int  # build.py:114: 40 for int3 else 32
    sym: bool = True  # --quantization-sym
    storage_nbit: int = 32  # --quantization-storage-nbit

    @property
    def nbit(self) -> int:  # quantization.py:313
        return int(self.mode[-1])

    @property
    def npu(self) -> int:
        return self.storage_nbit // self.nbit

    @property
    def maxint(self) -> int:
        return (1 << (self.nbit - 1)) - 1

    @property
    def fmt_id(self) -> int:
        return FMT_IDS[self.name]

    @property
    def torch_dtype(self) -> torch.dtype:
        return torch.float16 if self.dtype == "float16" else torch.float32

    def n_groups(self, K: int) -> int:
        return -(-K // self.group_size)

    def n_words(self, K: int) -> int:
        return -(-self.group_size // self.npu) * self.n_groups(K)

    @property
    def packed_torch_dtype(self) -> torch.dtype:
        """Storage words as torch sees them (torch has no unsigned 16 / 32-bit arithmetic types: same bits)."""
        return torch.int16 if self.storage_nbit == 16 else torch.int32

    @property
    def is_variant(self) -> bool:
        """uint16-storage int3 / fp4: encode, dequantize, take_decode and single-op GEMV only."""
        return self.storage_nbit != 32 or self.mode == "fp4"

    @property
    def has_min(self) -> bool:
        """A separate group-minimum tensor (the f16 asymmetric scheme; the f32 one packs bf16 scale | bf16 min)."""
        return (not self.sym) and self.dtype == "float16"

    @property
    def scale_torch_dtype(self) -> torch.dtype:
        return torch.int32 if (not self.sym and self.dtype == "float32") else self.torch_dtype

    def gemv_bytes(self, K: int, N: int, epilogue: str = "none") -> int:
        """Algorithmic HBM bytes of one GEMV."""
        sb = 2 if self.dtype == "float16" else 4
        ob = 4 if (epilogue == "cast_f32" or self.dtype == "float32") else 2
        b = (self.storage_nbit // 8) * self.n_words(K) * N + sb * self.n_groups(K) * N * (2 if self.has_min else 1) + sb * K + ob * N
        if epilogue in ("add", "multiply"):
            b += ob * N
        if epilogue in ("bias", "bias_gelu"):
            b += sb * N
        return b


SPECS = {
    "q4f16_0": QuantSpec("q4f16_0", "int4", "float16", 32),
    "q3f16_0": QuantSpec("q3f16_0", "int3", "float16", 40),
    "q4f32_0": QuantSpec("q4f32_0", "int4", "float32", 32),
    "q8f16_0": QuantSpec("q8f16_0", "int8", "float16", 32),
    # the reference CLI's default: asymmetric int4 (build.py:19 --quantization-sym off; quantization.py:105-140,183-211)
    "q4f16_asym": QuantSpec("q4f16_asym", "int4", "float16", 32, sym=False),
    "q4f32_asym": QuantSpec("q4f32_asym", "int4", "float32", 32, sym=False),
    # the remaining flag combinations: int3 in uint16 words (ios/README.md:39), fp4
    "q3f16_u16": QuantSpec("q3f16_u16", "int3", "float16", 40, storage_nbit=16),
    "fp4f16_0": QuantSpec("fp4f16_0", "fp4", "float16", 32),
    "fp4f32_0": QuantSpec("fp4f32_0", "fp4", "float32", 32),
}


def spec_from_flags(dtype: str, mode: str, sym: bool, storage_nbit: int, group_size: int | None = None) -> QuantSpec:
    """The reference's CLI flags -> QuantSpec (sym=False is the reference's default: the asymmetric int4 schemes).
    Raises for the combinations that are not on this path (uint16/uint8 storage, fp4: SURVEY.md 8(f) rank 4)."""
    if storage_nbit == 8:
        raise NotImplementedError("uint8 storage: no kernel of the reference uses it (build.py:22-25 lists it, dispatch/ has none)")
    if not sym and mode != "int4":
        raise NotImplementedError("asymmetric quantization exists for int4 only (quantization.py:283-284 asserts sym for int3 / fp4)")
    for s in SPECS.values():
        if s.mode == mode and s.dtype == dtype and s.sym == sym and s.storage_nbit == (32 if mode == "fp4" else storage_nbit):
            if group_size is not None and group_size != s.group_size:
                raise NotImplementedError(f"group size {group_size} (the reference hard-wires {s.group_size})")
            return s
    raise NotImplementedError(f"{mode}/{dtype}")
