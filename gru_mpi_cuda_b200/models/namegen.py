"""NameGenerator — the public model API.

Mirrors the reference's three entry points (`/root/reference/namegensf.cu`):

    namegen_initialize(N, rng_seed, parameter_fname)   :359  -> NameGenerator(...)
    namegen(N, random_floats, output)                  :627  -> .generate(N, random_floats)
    namegen_finalize()                                 :897  -> .finalize()

`generate` keeps the reference's contract: `random_floats` is `N*MAX_LEN` host
floats significant on the root rank, the result is `N x (MAX_LEN+1)` bytes on
the root rank, names are split over ranks in contiguous blocks of `N // size`
and the `N % size` tail stays zero (`:628-630`, `:640`).
"""
from __future__ import annotations

import ctypes
from pathlib import Path

import torch

from .. import _native
from ..config import GRUConfig
from ..parallel.partition import partition
from ..utils import weights as W

MODE_EXACT = 0
MODE_TC = 1
_MODES = {"exact": MODE_EXACT, "fp32": MODE_EXACT, "tc": MODE_TC, "bf16": MODE_TC}


class NameGenerator:
    """One rank's model instance on one GPU.

    Parameters
    ----------
    cfg : GRUConfig
    params : flat fp32 tensor in file order, or a path to the weight file
    device : CUDA device index (default: current device).  The reference never
        binds a device (`cudaSetDevice` is absent); here it is explicit.
    max_batch : most names this rank will run in one `generate` call
    mode : "tc"/"bf16" (tcgen05 tensor cores) or "exact"/"fp32" (bit-identical
        to the CPU oracle)
    comm : optional `parallel.dp.DPContext` for multi-GPU runs
    """

    def __init__(self, cfg: GRUConfig, params, device: int | None = None, max_batch: int = 4096,
                 mode: str = "tc", comm=None, tp_rank: int = 0, tp_world: int = 1):
        if not torch.cuda.is_available():
            raise RuntimeError("NameGenerator needs a CUDA device (use models.oracle / TorchNameGen on CPU)")
        self.cfg = cfg
        self.mode = _MODES[mode]
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.max_batch = int(max_batch)
        self.comm = comm
        if isinstance(params, (str, Path)):
            params = W.load(params, cfg)
        params = params.detach().to(torch.float32).contiguous().cpu()
        if params.numel() != W.param_count(cfg):
            raise ValueError(f"parameter vector has {params.numel()} floats, config needs {W.param_count(cfg)}")
        self._lib = _native.lib()
        handle = ctypes.c_void_p()
        c = cfg.to_c()
        self.tp_rank, self.tp_world = int(tp_rank), int(tp_world)
        _native.check(self._lib.b200gru_model_create_tp(ctypes.byref(c), params.data_ptr(), self.device,
                                                        self.max_batch, self.tp_rank, self.tp_world,
                                                        ctypes.byref(handle)), "model_create")
        self._h = handle
        dev = torch.device("cuda", self.device)
        # device-side staging for the single-process path (root == this rank)
        self._rf_dev = torch.empty(self.max_batch * cfg.max_len, dtype=torch.float32, device=dev)
        self._out_dev = torch.zeros(self.max_batch, cfg.max_len + 1, dtype=torch.uint8, device=dev)
        self._out_host = torch.zeros(self.max_batch, cfg.max_len + 1, dtype=torch.uint8).pin_memory()
        self.last_launches = 0
        self._copy_stream = None
        self._ready_ev = None

    # ------------------------------------------------------------------ core
    def generate_device(self, rfloats_dev: torch.Tensor, n: int, out_dev: torch.Tensor | None = None,
                        mode: str | None = None, done_flag_ptr: int = 0, done_value: int = 0,
                        ready_event: torch.cuda.Event | None = None) -> torch.Tensor:
        """Generate `n` names from device-resident floats; asynchronous on the
        current stream.  `rfloats_dev` / `out_dev` may live on a peer GPU."""
        cfg = self.cfg
        if n > self.max_batch:
            raise ValueError(f"n={n} > max_batch={self.max_batch}")
        if out_dev is None:
            out_dev = self._out_dev
        io = _native.CIo(rfloats_dev.data_ptr(), out_dev.data_ptr(), done_flag_ptr or None, done_value,
                         None, 0, ready_event.cuda_event if ready_event is not None else None)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        m = self.mode if mode is None else _MODES[mode]
        _native.check(self._lib.b200gru_generate(self._h, m, n, ctypes.byref(io), ctypes.c_void_p(stream)),
                      "generate")
        self.last_launches = self._lib.b200gru_last_launch_count(self._h)
        return out_dev[:n]

    def generate(self, n_names: int, random_floats: torch.Tensor, mode: str | None = None,
                 out: torch.Tensor | None = None) -> torch.Tensor:
        """namegen(): host floats in, host `uint8[n_names, max_len+1]` out.

        `out` plays the reference's caller-allocated `char *output`
        (`namegensf.cu:624-627`): a pinned host `uint8[n_names, max_len+1]` tensor
        that is filled and returned instead of a fresh copy."""
        if self.comm is not None:
            return self.comm.generate(self, n_names, random_floats, mode=mode, out=out)
        cfg = self.cfg
        need = n_names * cfg.max_len
        rf = random_floats.reshape(-1)
        if rf.numel() < need:
            raise ValueError(f"random_floats has {rf.numel()} values, need {need}")
        sl = partition(n_names, 1, 0)
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream()
            # the floats go up on a side stream; the kernels only wait for them
            # (an event) right before the first sampling kernel
            if self._copy_stream is None:
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._ready_ev = torch.cuda.Event()
            self._copy_stream.wait_stream(main)
            with torch.cuda.stream(self._copy_stream):
                self._rf_dev[:need].copy_(rf[:need], non_blocking=True)
                self._ready_ev.record(self._copy_stream)
            rows = self.generate_device(self._rf_dev, sl.count, mode=mode, ready_event=self._ready_ev)
            host = self._out_host[:n_names] if out is None else out
            host.copy_(rows, non_blocking=True)
            main.synchronize()
        return host.clone() if out is None else host

    def enable_logits_tap(self, on: bool = True) -> None:
        """Make the tensor-core FC kernel store its logits too (slow; tests only)."""
        _native.check(self._lib.b200gru_model_set_debug(self._h, 1 if on else 0))

    def set_path(self, path: str = "auto", persist_max_names: int = -1) -> bool:
        """Kernel path of the tensor-core mode: "auto" (persistent kernel for
        calls of at most `persist_max_names` names, multi-launch above),
        "multi" or "persistent".  Returns whether the persistent kernel can run
        this model at all (its weight slices must fit in shared memory)."""
        code = {"auto": 0, "multi": 1, "persistent": 2}[path]
        return bool(_native.check(self._lib.b200gru_model_set_path(self._h, code, persist_max_names), "set_path"))

    def set_compaction(self, on: bool = True) -> None:
        """Multi-launch path: pack still-running names to the front of the batch
        a few times per call so finished names stop costing GEMM rows."""
        _native.check(self._lib.b200gru_model_set_compaction(self._h, 1 if on else 0), "set_compaction")

    def check(self) -> None:
        """Debug build (`B200GRU_DEBUG=1`): verify the guard bands around every
        device buffer of the model now; raises on damage.  No-op in release."""
        _native.check(self._lib.b200gru_model_check(self._h), "model_check")

    def set_trace(self, on: bool = True) -> None:
        """Record a CUDA event after every launch of the tensor-core path."""
        _native.check(self._lib.b200gru_model_set_trace(self._h, 1 if on else 0))

    def trace(self) -> list[tuple[str, float]]:
        """(kernel label, device ms) per launch of the last traced generate."""
        cap = 256
        ms = (ctypes.c_float * cap)()
        names = (ctypes.c_char_p * cap)()
        n = _native.check(self._lib.b200gru_model_get_trace(self._h, ms, names, cap), "get_trace")
        return [(names[i].decode(), ms[i]) for i in range(n)]

    def persist_timeline(self) -> dict:
        """In-kernel timeline of the last traced persistent generate:
        `{(cta, role): [(t_us, step, layer, event), ...]}` with cta in
        ("weights0", "fc0"), role in ("tma", "mma", "epilogue"), times relative to
        the first stamp.  Events are listed in include/b200gru.h."""
        words = 2 * 3 * 256
        buf = (ctypes.c_uint64 * words)()
        n = _native.check(self._lib.b200gru_model_persist_timeline(self._h, buf, words), "persist_timeline")
        vals = [v for v in buf[:n] if v]
        if not vals:
            return {}
        t0 = min(v >> 16 for v in vals)
        out = {}
        for ci, cta in enumerate(("weights0", "fc0")):
            for ri, role in enumerate(("tma", "mma", "epilogue")):
                base = (ci * 3 + ri) * 256
                ev = [((v >> 16) - t0) / 1e3 for v in buf[base:base + 256] if v]
                codes = [v & 0xffff for v in buf[base:base + 256] if v]
                out[(cta, role)] = [(t, c >> 8, (c >> 4) & 15, c & 15) for t, c in zip(ev, codes)]
        return out

    def debug_logits(self, n: int) -> torch.Tensor:
        """Last step's pre-softmax FC outputs for the first `n` names."""
        out = torch.empty(n, self.cfg.num_char, dtype=torch.float32, device=torch.device("cuda", self.device))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _native.check(self._lib.b200gru_debug_logits(self._h, out.data_ptr(), n, ctypes.c_void_p(stream)))
        return out

    @property
    def device_bytes(self) -> int:
        return self._lib.b200gru_model_device_bytes(self._h)

    def finalize(self) -> None:
        if getattr(self, "_h", None):
            _native.check(self._lib.b200gru_model_destroy(self._h), "model_destroy")
            self._h = None

    close = finalize

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.finalize()

    def __del__(self):
        try:
            self.finalize()
        except Exception:
            pass


def decode(rows: torch.Tensor, eos: int | None = None) -> list[bytes]:
    """Rows of the output matrix as byte strings, cut at the first NUL (the way
    a C caller would `printf("%s")` them) and optionally at EOS."""
    out = []
    for row in rows.cpu().numpy():
        b = bytes(row.tolist())
        b = b.split(b"\0", 1)[0]
        if eos is not None and eos != 0:
            b = b.split(bytes([eos]), 1)[0]
        out.append(b)
    return out
