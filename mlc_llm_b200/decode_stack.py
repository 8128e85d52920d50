"""Layer-sequence driver for the decode step: the quantized-Linear launches of one token, in the
reference's order, standing in for the compiled Relax `decoding` function
(mlc_llm/relax_model/llama.py:581-611 built from LlamaDecoderLayer :340-401; GPT-NeoX:
mlc_llm/relax_model/gpt_neox.py:134-288; call stack SURVEY.md 3.2).

Per Llama layer the reference launches 7 fused dequant-GEMVs (q, k, v, o+residual, gate+silu,
up*gate, down+residual) and one more for lm_head (cast to f32): 225 per token for 7B.  Here
q|k|v share one launch (they read the same activation), everything is captured in one CUDA graph
with programmatic dependent launch, and - for tensor parallelism - q/k/v/gate/up are
column-sharded, o/down row-sharded, with one all-reduce per sharded block.

Attention, norms and rotary embedding are NOT part of this path: the
activation that would come out of attention is taken from the q projection so that the
dependency chain between launches is the real one.

Weights are synthetic: W ~ N(0, 0.02^2) (HF initializer_range, llama.py:24; residual-branch output
projections 0.02 / sqrt(2 L), see _build_weights) drawn on the device
and packed by the library's own encoder (b200q_encode, bit-exact against the reference's
te_encode_sym - tests/test_gpu_parity.py).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Tuple

import os

import torch

from . import _lib, ops
from .quant_spec import SPECS, QuantSpec


@dataclasses.dataclass(frozen=True)
class ModelShape:
    name: str
    family: str  # "llama" | "gpt_neox"
    hidden: int
    inter: int
    layers: int
    vocab: int


# llama.py:18-22,51-54; gpt_neox.py:50-57,99-107; 13B / 65B are BASELINE.json shapes
MODEL_SHAPES = {
    "llama-7b": ModelShape("llama-7b", "llama", 4096, 11008, 32, 32000),
    "llama-13b": ModelShape("llama-13b", "llama", 5120, 13824, 40, 32000),
    "llama-65b": ModelShape("llama-65b", "llama", 8192, 22016, 80, 32000),
    "redpajama-3b": ModelShape("redpajama-3b", "gpt_neox", 2560, 10240, 32, 50432),
    "dolly-v2-3b": ModelShape("dolly-v2-3b", "gpt_neox", 2560, 10240, 32, 50280),
}


def linear_list(shape: ModelShape) -> List[Tuple[str, int, int, str]]:
    """(name, K, N, epilogue) of every quantized Linear of ONE layer, in launch order."""
    h, i = shape.hidden, shape.inter
    if shape.family == "llama":
        return [("q", h, h, "none"), ("k", h, h, "none"), ("v", h, h, "none"), ("o", h, h, "add"),
                ("gate", h, i, "silu"), ("up", h, i, "multiply"), ("down", i, h, "add")]
    return [("q", h, h, "bias"), ("k", h, h, "bias"), ("v", h, h, "bias"), ("dense", h, h, "bias"),
            ("h_to_4h", h, i, "bias_gelu"), ("4h_to_h", i, h, "bias")]


def token_bytes(shape: ModelShape, spec: QuantSpec, tp: int = 1) -> int:
    """Algorithmic HBM bytes of one token's quantized-Linear launches on ONE rank."""
    total = 0
    for name, K, N, ep in linear_list(shape):
        if name in ("o", "down", "dense", "4h_to_h"):
            K //= tp
        else:
            N //= tp
        total += spec.gemv_bytes(K, N, ep)
    total *= shape.layers
    total += spec.gemv_bytes(shape.hidden, shape.vocab // tp, "cast_f32")
    return total


def token_macs(shape: ModelShape) -> int:
    return shape.layers * sum(K * N for _, K, N, _ in linear_list(shape)) + shape.hidden * shape.vocab


class QuantWeight:
    """One packed Linear weight: packed uint32[n_words(K), N] + scale dtype[n_groups(K), N]."""

    def __init__(self, packed: torch.Tensor, scale: torch.Tensor, K: int, N: int):
        self.packed, self.scale, self.K, self.N = packed, scale, K, N


def shard_packed(packed, scale, spec: QuantSpec, K: int, N: int, col_slice=None, row_slice=None):
    """Tensor-parallel shard of one packed Linear: pure indexing, no repacking.

    col_slice=(r, tp): output columns [r*N/tp, (r+1)*N/tp)  - packed[:, n0:n1], scale[:, n0:n1]
    row_slice=(r, tp): input rows     [r*K/tp, (r+1)*K/tp)  - whole groups: packed word rows / scale rows
    Works on any torch tensor (CPU tensors in the gloo tests).  Returns (packed, scale, K, N) of the shard."""
    if col_slice is not None:
        r, tp = col_slice
        n = N // tp
        if n % 8 or N % tp:
            raise ValueError(f"column shard {N}/{tp} is not a multiple of 8")
        packed = packed[:, r * n:(r + 1) * n].contiguous()
        scale = scale[:, r * n:(r + 1) * n].contiguous()
        N = n
    if row_slice is not None:
        r, tp = row_slice
        k = K // tp
        if K % tp or k % spec.group_size:
            raise ValueError(f"row shard {K}/{tp} is not a multiple of the group size {spec.group_size}")
        wpg = spec.n_words(spec.group_size)
        gk = k // spec.group_size
        packed = packed[r * gk * wpg:(r + 1) * gk * wpg].contiguous()
        scale = scale[r * gk:(r + 1) * gk].contiguous()
        K = k
    return packed, scale, K, N


def synth_weight(spec: QuantSpec, K: int, N: int, seed: int, device, col_slice=None, row_slice=None,
                 std: float = 0.02) -> QuantWeight:
    """Draw W[N, K] ~ N(0, std^2) on the device, pack it with the library encoder, slice for TP.

    col_slice=(r, tp): keep output columns [r*N/tp, (r+1)*N/tp)     (column parallel)
    row_slice=(r, tp): keep input rows    [r*K/tp, (r+1)*K/tp)      (row parallel; K/tp % group == 0)"""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    W = torch.empty((N, K), dtype=torch.float32, device=device)
    W.normal_(0.0, std, generator=g)
    W = W.to(spec.torch_dtype)
    packed, scale = ops.encode(W, spec, transpose=True)
    del W
    packed, scale, K, N = shard_packed(packed, scale, spec, K, N, col_slice, row_slice)
    return QuantWeight(packed, scale, K, N)


class DecodeStack:
    """The quantized-Linear stream of one decode step, graph-captured.

    decode_step(): hidden state already on the device (x_in) -> logits on the device.
    decode_step_host(x_host, logits_host): pinned host buffers in and out (what the reference's
    LLMChatModule::Forward does around `decoding`, cpp/llm_chat.cc:920-946: token in, logits out).
    """

    def __init__(self, shape, spec, device=None, tp_rank: int = 0, tp_size: int = 1, seed: int = 0xB200,
                 layers: Optional[int] = None, process_group=None, fuse_qkv: bool = True, fuse_gate_up: bool = True,
                 pdl: bool = True, use_graph: bool = True, use_program: bool = False, segments: str = "fused",
                 weights_from: Optional["DecodeStack"] = None):
        self.shape = MODEL_SHAPES[shape] if isinstance(shape, str) else shape
        self.spec = SPECS[spec] if isinstance(spec, str) else spec
        self.device = torch.device(device if device is not None else "cuda")
        self.tp_rank, self.tp_size, self.pg = tp_rank, tp_size, process_group
        self.n_layers = self.shape.layers if layers is None else layers
        self.fuse_qkv, self.fuse_gate_up, self.pdl, self.use_graph = fuse_qkv, fuse_gate_up, pdl, use_graph
        if tp_size > 1 and self.spec.dtype != "float16":
            raise NotImplementedError("tensor parallel decode is wired for the f16 schemes")
        self.dt = self.spec.torch_dtype
        # segments: where a program launch ends.  "fused": only at collectives (the GEMV chain of a whole token is one
        # launch).  "reference": additionally wherever the reference's `decoding` function has a non-GEMV call_tir
        # between two fused dequant-GEMVs (attention after q|k|v, the norm after o and after down;
        # mlc_llm/relax_model/llama.py:340-401) - what a drop-in behind the reference's operator boundary can do
        # today, because those foreign kernels read and write plain buffers, not the program's tagged words.
        if segments not in ("fused", "reference"):
            raise ValueError("segments: 'fused' or 'reference'")
        self.segments = segments
        if weights_from is not None:  # share the packed tensors of another stack (same shape / scheme / sharding)
            o = weights_from
            if (o.shape, o.spec, o.tp_rank, o.tp_size, o.n_layers) != (self.shape, self.spec, tp_rank, tp_size, self.n_layers):
                raise ValueError("weights_from: stacks differ")
            self.layers, self.biases, self.lm_head = o.layers, o.biases, o.lm_head
        else:
            self._build_weights(seed)
        self._build_buffers()
        self.graph = None
        self.launches_per_step = 0
        self.collectives_per_step = 0
        # use_program: the whole GEMV chain as persistent program launches (one per all-reduce-free segment)
        self.use_program = use_program
        # tensor parallel + program + q4f16_0: the per-block all-reduce is fused into the program (one launch per
        # token and rank; only the final logits all-gather is an NCCL call)
        self.fused_tp = bool(use_program and tp_size > 1 and ops.program_uses_tensor_cores(self.spec) and
                             os.environ.get("B200Q_FUSED_TP", "1") != "0")
        self._plan = None
        self._steps = None
        # launch trace for snapshot(): which Linear(s) every launch runs on which buffers, and how every buffer is
        # laid out over the tensor-parallel ranks
        self._trace: Optional[list] = None
        self._layout: Dict[int, str] = {}
        self._row_ep: Dict[int, tuple] = {}
        if use_program:
            self._build_plan()

    # ------------------------------------------------------------------ weights
    def _build_weights(self, seed: int):
        sh, sp, tp, r = self.shape, self.spec, self.tp_size, self.tp_rank
        self.layers: List[Dict[str, QuantWeight]] = []
        self.biases: List[Dict[str, torch.Tensor]] = []
        row_parallel = ("o", "down", "dense", "4h_to_h")
        # HF initializer_range 0.02 (llama.py:24); the projections that write into the residual stream get the usual
        # depth scaling 0.02 / sqrt(2 L) (GPT-2 / GPT-NeoX output-layer init): this stack has no norm layers (they are
        # not on the path), and without it the un-normalized residual stream overflows f16 after a few layers
        std_out = 0.02 / (2.0 * sh.layers) ** 0.5
        for li in range(self.n_layers):
            ws, bs = {}, {}
            for j, (name, K, N, ep) in enumerate(linear_list(sh)):
                s = seed * 1000003 + li * 16 + j
                std = std_out if name in row_parallel else 0.02
                if tp > 1 and name in row_parallel:
                    ws[name] = synth_weight(sp, K, N, s, self.device, row_slice=(r, tp), std=std)
                elif tp > 1:
                    ws[name] = synth_weight(sp, K, N, s, self.device, col_slice=(r, tp), std=std)
                else:
                    ws[name] = synth_weight(sp, K, N, s, self.device, std=std)
                if ep.startswith("bias"):
                    g = torch.Generator(device=self.device)
                    g.manual_seed(s + 7)
                    b = torch.empty(N, dtype=torch.float32, device=self.device).normal_(0.0, 0.02, generator=g).to(self.dt)
                    if tp > 1 and name not in row_parallel:
                        n = N // tp
                        b = b[r * n:(r + 1) * n].contiguous()
                    bs[name] = b
            self.layers.append(ws)
            self.biases.append(bs)
        self.lm_head = synth_weight(sp, sh.hidden, sh.vocab, seed * 1000003 + 999983, self.device,
                                    col_slice=(r, tp) if tp > 1 else None)

    def weight_bytes(self) -> int:
        tot = 0
        for ws in self.layers:
            for w in ws.values():
                tot += w.packed.numel() * 4 + w.scale.numel() * w.scale.element_size()
        return tot + self.lm_head.packed.numel() * 4 + self.lm_head.scale.numel() * self.lm_head.scale.element_size()

    # ------------------------------------------------------------------ activations
    def _build_buffers(self):
        sh, tp, dev, dt = self.shape, self.tp_size, self.device, self.dt
        h, i = sh.hidden, sh.inter
        e = lambda n, d=dt: torch.zeros(n, dtype=d, device=dev)
        self.x_in = e(h)
        self.buf = []
        for _ in range(self.n_layers):
            self.buf.append({"q": e(h // tp), "k": e(h // tp), "v": e(h // tp), "x1": e(h), "g": e(i // tp),
                             "u": e(i // tp), "a": e(i // tp), "x2": e(h)})
        # split-K scratch of the GEMV kernels: owned by this stack (one stream), zeroed once
        self.ws = torch.zeros(int(_lib.lib().b200q_gemv_workspace_bytes()), dtype=torch.uint8, device=dev)
        self.logits_part = e(sh.vocab // tp, torch.float32)
        self.logits = self.logits_part if tp == 1 else e(sh.vocab, torch.float32)

    # ------------------------------------------------------------------ the launch sequence
    def _launch(self, items, x, x_mul=None, reduce=False, row=False):
        """One kernel launch: items = [(weight, epilogue, ep0, out), ...] all reading x (* x_mul)."""
        if self._trace is not None:
            self._trace.append(dict(x=x, x_mul=x_mul, row=row,
                                    ops=[dict(w=w, epilogue=ep, ep0=e0, out=o) for w, ep, e0, o in items]))
            for _, _, _, o in items:
                if self.tp_size > 1:
                    self._layout[o.data_ptr()] = "rep" if row else "col"  # (reduced outputs are complete on every rank)
            if self._trace_only:
                return
        if self._steps is not None:  # recording a program
            self._steps.append(dict(weights=[(w.packed, w.scale) for w, _, _, _ in items], x=x, x_mul=x_mul,
                                    epilogues=[ep for _, ep, _, _ in items], ep0s=[e0 for _, _, e0, _ in items],
                                    outs=[o for _, _, _, o in items], reduces=[reduce] * len(items)))
            return
        ops.gemv_multi([(w.packed, w.scale) for w, _, _, _ in items], x, self.spec,
                       epilogues=[ep for _, ep, _, _ in items], ep0s=[e0 for _, _, e0, _ in items],
                       outs=[o for _, _, _, o in items], pdl=self.pdl, x_mul=x_mul, ws=self.ws)
        self.launches_per_step += 1

    def _gemv(self, w: QuantWeight, x, ep, ep0, out):
        self._launch([(w, ep, ep0, out)], x)

    def _flush_steps(self):
        if self._steps:
            tp = (self.tp_rank, self.tp_size, self.pg) if self.fused_tp else None
            self._plan.append(("program", ops.GemvProgram(self._steps, self.spec, tp=tp)))
        self._steps = []

    def _segment_break(self):
        """A call site of the reference's `decoding` where a non-GEMV kernel sits between two fused GEMVs."""
        if self.segments == "reference" and self._steps is not None and not self._trace_only:
            self._flush_steps()

    def _build_plan(self):
        """Record the step's launches once: GEMV programs separated by the collectives."""
        self._plan, self._steps = [], []
        self._run_sequence()
        self._flush_steps()
        self._steps = None

    def _run_plan(self):
        import torch.distributed as dist

        self.launches_per_step = 0
        self.collectives_per_step = 0
        for kind, obj in self._plan:
            if kind == "program":
                obj.run()
                self.launches_per_step += 1
            elif kind == "all_reduce":
                dist.all_reduce(obj, group=self.pg)
                self.collectives_per_step += 1
            else:
                dist.all_gather_into_tensor(obj[0], obj[1], group=self.pg)
                self.collectives_per_step += 1
        return self.logits

    def _all_reduce(self, t):
        if self._trace is not None and self._trace_only:
            return
        if self._steps is not None:
            self._flush_steps()
            self._plan.append(("all_reduce", t))
            return
        import torch.distributed as dist

        dist.all_reduce(t, group=self.pg)
        self.collectives_per_step += 1

    def _row_parallel(self, w, x, residual, out, ep_single, x_mul=None):
        """o / down: out = residual + sum_ranks(x_r @ w_r).  Rank 0 carries the residual (or bias)
        into the sum so the reduction result is the final value."""
        if self.fused_tp and (self._steps is not None or self._trace_only):
            # the reduction over the ranks happens inside the program launch (peer pushes over NVLink): the
            # epilogue is applied once, by the rank that owns the output panel, to the f32 sum of the partials
            self._launch([(w, ep_single, residual, out)], x, x_mul, reduce=True, row=True)
            return
        self._row_ep[out.data_ptr()] = (ep_single, residual)
        if self.tp_size == 1 or self.tp_rank == 0:
            self._launch([(w, ep_single, residual, out)], x, x_mul, row=True)
        else:
            self._launch([(w, "none", None, out)], x, x_mul, row=True)
        if self.tp_size > 1:
            self._all_reduce(out)

    def _run_llama(self):
        x = self.x_in
        for li in range(self.n_layers):
            w, b = self.layers[li], self.buf[li]
            if self.fuse_qkv:
                self._launch([(w[n], "none", None, b[n]) for n in ("q", "k", "v")], x)
            else:
                for n in ("q", "k", "v"):
                    self._gemv(w[n], x, "none", None, b[n])
            self._segment_break()  # rotary + KV append + attention + softmax (llama.py:233-290)
            attn = b["q"]  # stand-in for the attention output (same shape, same dependency)
            self._row_parallel(w["o"], attn, x, b["x1"], "add")
            self._segment_break()  # post_attention_layernorm (llama.py:387)
            if self.fuse_gate_up:
                # gate|up in one launch; the SwiGLU product silu(gate) * up is formed while down_proj
                # stages its activation (bit-identical to the reference's `multiply` epilogue)
                self._launch([(w["gate"], "silu", None, b["g"]), (w["up"], "none", None, b["u"])], b["x1"])
                self._row_parallel(w["down"], b["g"], b["x1"], b["x2"], "add", x_mul=b["u"])
            else:
                self._gemv(w["gate"], b["x1"], "silu", None, b["g"])
                self._gemv(w["up"], b["x1"], "multiply", b["g"], b["a"])
                self._row_parallel(w["down"], b["a"], b["x1"], b["x2"], "add")
            self._segment_break()  # the next layer's input_layernorm / the final norm (llama.py:361, 523)
            x = b["x2"]
        return x

    def _run_neox(self):
        x = self.x_in
        for li in range(self.n_layers):
            w, b, bias = self.layers[li], self.buf[li], self.biases[li]
            if self.fuse_qkv:
                self._launch([(w[n], "bias", bias[n], b[n]) for n in ("q", "k", "v")], x)
            else:
                for n in ("q", "k", "v"):
                    self._gemv(w[n], x, "bias", bias[n], b[n])
            self._segment_break()  # attention (gpt_neox.py:134-200)
            self._row_parallel(w["dense"], b["q"], bias["dense"], b["x1"], "bias")
            self._segment_break()  # post_attention_layernorm
            self._gemv(w["h_to_4h"], b["x1"], "bias_gelu", bias["h_to_4h"], b["a"])
            self._row_parallel(w["4h_to_h"], b["a"], bias["4h_to_h"], b["x2"], "bias")
            self._segment_break()  # next layer's input_layernorm / final_layer_norm
            x = b["x2"]
        return x

    def _run_sequence(self):
        x = self._run_llama() if self.shape.family == "llama" else self._run_neox()
        self._gemv(self.lm_head, x, "none" if self.spec.dtype == "float32" else "cast_f32", None, self.logits_part)
        if self.tp_size > 1:
            if self._trace is not None and self._trace_only:
                return self.logits
            if self._steps is not None:
                self._flush_steps()
                self._plan.append(("all_gather", (self.logits, self.logits_part)))
                return self.logits
            import torch.distributed as dist

            dist.all_gather_into_tensor(self.logits, self.logits_part, group=self.pg)
            self.collectives_per_step += 1
        return self.logits

    def _run(self):
        if self._plan is not None:
            return self._run_plan()
        self.launches_per_step = 0
        self.collectives_per_step = 0
        return self._run_sequence()

    # ------------------------------------------------------------------ verification support
    _trace_only = False

    def _record_trace(self):
        """Walk the launch sequence once without launching anything: which Linear runs on which buffers."""
        steps, plan = self._steps, self._plan
        self._steps, self._plan = None, None
        self._trace, self._trace_only = [], True
        try:
            self._run_sequence()
        finally:
            self._trace_only = False
            self._steps, self._plan = steps, plan
        trace, self._trace = self._trace, None
        return trace

    def _weight_name(self, w: "QuantWeight"):
        for li, ws in enumerate(self.layers):
            for name, cand in ws.items():
                if cand is w:
                    return li, name
        return self.n_layers, "lm_head"

    def _full_host(self, t: Optional[torch.Tensor]):
        """The complete (unsharded) tensor as a host array: column-parallel outputs are concatenated over the
        ranks; reduced (row-parallel) outputs and the input are complete on every rank."""
        if t is None:
            return None
        kind = self._layout.get(t.data_ptr(), "rep") if self.tp_size > 1 else "rep"
        if kind == "rep":
            return t.detach().cpu().numpy().copy()
        import torch.distributed as dist

        parts = [torch.empty_like(t) for _ in range(self.tp_size)]
        dist.all_gather(parts, t.contiguous(), group=self.pg)
        return torch.cat([p.reshape(-1) for p in parts]).cpu().numpy()  # "col"

    def snapshot(self):
        """After a decode_step(): for every launch of the step, the host copies of the FULL activation it read
        and of the FULL outputs it wrote (every rank must call this - tensor-parallel shards are gathered; the
        checker, which is not part of this package, compares them operator by operator with the CPU oracle)."""
        torch.cuda.synchronize(self.device)
        trace = self._record_trace()
        out = []
        for ln in trace:
            x = self._full_host(ln["x"])
            x_mul = self._full_host(ln["x_mul"])
            ops_ = []
            for op in ln["ops"]:
                li, name = self._weight_name(op["w"])
                ep = op["epilogue"]
                ep0 = op["ep0"]
                if self.tp_size > 1 and ln["row"] and not self.fused_tp:
                    # NCCL chain: rank 0 carries the residual / bias into the all-reduced sum
                    ep, ep0 = self._row_ep.get(op["out"].data_ptr(), (ep, ep0))
                ops_.append(dict(layer=li, name=name, epilogue=ep, ep0=self._full_host(ep0), out=self._full_host(op["out"])))
            out.append(dict(x=x, x_mul=x_mul, ops=ops_))
        return out

    def capture(self):
        """Warm up on a side stream and capture the step into one CUDA graph."""
        if not self.use_graph:
            return
        st = torch.cuda.Stream(device=self.device)
        st.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(st):
            for _ in range(2):
                self._run()
            st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                self._run()
        torch.cuda.current_stream(self.device).wait_stream(st)
        self.graph = g

    def decode_step(self) -> torch.Tensor:
        if self.graph is not None:
            self.graph.replay()
        else:
            self._run()
        return self.logits

    def decode_step_host(self, x_host: torch.Tensor, logits_host: torch.Tensor) -> torch.Tensor:
        self.x_in.copy_(x_host, non_blocking=True)
        self.decode_step()
        logits_host.copy_(self.logits, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return logits_host
