"""ZeRO-1 optimizers on flat buffers.

Reference: ``src/neuronx_distributed/optimizer/zero_redundancy_optimizer.py:29-362`` -- thin subclasses of
torch_xla's ``ZeroRedundancyOptimizer`` (external native dependency) which shards every
parameter on dim 0 over the DP group, reduce-scatters grads, steps an inner optimizer on fp32 shards and
all-gathers the parameters.  torch_xla does not exist here, so the whole engine is ours.

B200 design
  * All local parameters of one (param-group, dtype) live in ONE contiguous *flat parameter buffer* and
    their grads in ONE *flat grad buffer* (``p.data`` / ``p.grad`` are views).  Each DP rank owns a
    contiguous ``1/dp`` slice of the flat buffer: fp32 master weights, ``exp_avg``, ``exp_avg_sq``.
  * ``step()`` = (dp>1: one reduce-scatter of the flat grad buffer) -> sum-of-squares of the local slice
    -> ONE small all-reduce of the scalar over the shard/TP/PP groups -> clip coefficient computed on the
    device (no host sync) -> ONE fused AdamW kernel over the slice that also writes the rounded bf16
    parameters back into the flat buffer -> (dp>1: one all-gather of the flat parameter buffer).
    With dp == 1 (pure TP) the step is three kernel launches for the whole model.
  * Buckets are ordered [TP-sharded params | sequence-parallel replicated params | other replicated
    params]: the replicated tail is counted ``1/tp`` in the global norm (reference ``grads.py:156-165``)
    and the SP region's grad all-reduce over TP (``grads.py:313-329``, K15) is ONE collective on a
    contiguous range.
  * Optimizers other than AdamW/Adam run through a generic path: the inner ``optimizer_class`` steps fp32
    master-slice parameters (same semantics as the reference, more launches).

Checkpoint state has the reference's top-level keys (``state``, ``base_state``, ``shape_info``,
``param_groups``, reference ``:284-303``); shards are 1-D slices of the flat layout (``layout`` key), which
``optimizer/convert_zero_checkpoints.py`` converts to/from full tensors or another DP size.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import optim as optim_ops
from ..parallel import parallel_state as ps
from ..utils.logger import get_logger

logger = get_logger()

_ALIGN = 128  # elements; keeps every shard boundary 256-byte aligned for bf16


def _group_size(g) -> int:
    return 1 if g is None else dist.get_world_size(group=g)


def _group_rank(g) -> int:
    return 0 if g is None else dist.get_rank(group=g)


class _Bucket:
    """One flat (param, grad) buffer pair and this rank's optimizer-state slice of it."""

    def __init__(self, params: List[torch.nn.Parameter], group_idx: int, shard_group, optimizer_dtype: torch.dtype,
                 grad_dtype: Optional[torch.dtype], symm=None, name: str = "zero1"):
        self.group_idx = group_idx
        self.symm = None
        self.param_ptrs = self.grad_ptrs = None
        self.shard_group = shard_group
        self.dp = _group_size(shard_group)
        self.dp_rank = _group_rank(shard_group)
        # order: TP-sharded | SP-tagged replicated | other replicated
        def cls(p):
            if getattr(p, "tensor_model_parallel", False) or getattr(p, "expert_model_parallel", False):
                return 0
            return 1 if getattr(p, "sequence_parallel_enabled", False) else 2
        self.params = sorted(params, key=cls)  # stable
        classes = [cls(p) for p in self.params]
        self.dtype = self.params[0].dtype
        self.device = self.params[0].device
        self.offsets: List[int] = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += ((p.numel() + 7) // 8) * 8  # keep every parameter 16-byte aligned (TMA operands)
        self.sp_begin = next((self.offsets[i] for i, c in enumerate(classes) if c >= 1), off)
        self.dup_other_begin = next((self.offsets[i] for i, c in enumerate(classes) if c >= 2), off)
        self.used = off
        quantum = self.dp * _ALIGN
        self.total = ((off + quantum - 1) // quantum) * quantum
        self.shard_size = self.total // self.dp
        self.lo = self.dp_rank * self.shard_size
        self.hi = self.lo + self.shard_size

        gdt = grad_dtype or self.dtype
        if (symm is not None and self.dp > 1 and self.dp <= 8 and self.device.type == "cuda"
                and self.dtype == torch.bfloat16 and gdt in (torch.bfloat16, torch.float32)
                and optimizer_dtype == torch.float32):
            # fused step (csrc/fused_zero1.cu): both flat buffers live in peer-mapped memory of the shard group
            import ctypes as _c

            esz = {torch.bfloat16: 2, torch.float32: 4}
            lp, ptab = symm.region(f"{name}:param", self.total * esz[self.dtype])
            lg, gtab = symm.region(f"{name}:grad", self.total * esz[gdt])
            self.flat_param = symm.tensor_view(lp, (self.total,), self.dtype)
            self.flat_grad = symm.tensor_view(lg, (self.total,), gdt)
            self.flat_param.zero_()
            self.flat_grad.zero_()
            self.param_ptrs = (_c.c_void_p * self.dp)(*[int(v) for v in ptab.tolist()])
            self.grad_ptrs = (_c.c_void_p * self.dp)(*[int(v) for v in gtab.tolist()])
            self.symm = symm
        else:
            self.flat_param = torch.zeros(self.total, dtype=self.dtype, device=self.device)
            self.flat_grad = torch.zeros(self.total, dtype=gdt, device=self.device)
        for p, o in zip(self.params, self.offsets):
            n = p.numel()
            view = self.flat_param[o:o + n].view(p.shape)
            with torch.no_grad():
                view.copy_(p.data)
            p.data = view
            gview = self.flat_grad[o:o + n].view(p.shape)
            if gdt == p.dtype:
                p.grad = gview
                p._nxd_flat_grad = gview   # wgrad GEMMs write here directly when ``p.grad is None`` (layers.py)
            else:
                p.main_grad = gview  # fp32 accumulation target (GEMM epilogue / hook), reference use_grad_acc_hook
        self.master = self.flat_param[self.lo:self.hi].to(optimizer_dtype).clone()
        self.exp_avg: Optional[torch.Tensor] = None
        self.exp_avg_sq: Optional[torch.Tensor] = None
        self.grad_shard = self.flat_grad[self.lo:self.hi] if self.dp == 1 else torch.zeros(
            self.shard_size, dtype=torch.float32 if self.symm is not None else gdt, device=self.device)

    def init_state(self, optimizer_dtype):
        if self.exp_avg is None:
            self.exp_avg = torch.zeros(self.shard_size, dtype=optimizer_dtype, device=self.device)
            self.exp_avg_sq = torch.zeros(self.shard_size, dtype=optimizer_dtype, device=self.device)

    def local_range(self, begin: int, end: int):
        """Intersection of [begin, end) with my shard, relative to the shard start."""
        a, b = max(begin, self.lo), min(end, self.hi)
        return (a - self.lo, b - self.lo) if b > a else (0, 0)


NO_SHARD_GROUP = "no-shard-group"


class NeuronZero1Optimizer(torch.optim.Optimizer):
    """ZeRO-1 over the DP group.  Constructor keywords follow the reference wiring
    (``trainer/trainer.py:244-269``)."""

    def __init__(self, params, optimizer_class=torch.optim.AdamW, optimizer_dtype: Optional[torch.dtype] = None,
                 pin_layout: bool = False, sharding_groups=None, grad_norm_groups=None, grad_clipping: bool = True,
                 max_norm: Optional[float] = 1.0, coalesce_cc: bool = True, use_grad_acc_hook: bool = False,
                 higher_cc_precision: bool = False, save_master_weights: bool = False, lazy_init: bool = False,
                 shard_process_group=None, **defaults):
        self.optimizer_class = optimizer_class
        self.optimizer_dtype = optimizer_dtype or torch.float32
        self.grad_clipping = grad_clipping
        self.max_norm = max_norm if max_norm is not None else 1.0
        self.use_grad_acc_hook = use_grad_acc_hook
        self.higher_cc_precision = higher_cc_precision
        self.save_master_weights = save_master_weights
        self.coalesce_cc = coalesce_cc
        self.pin_layout = pin_layout
        # ``NO_SHARD_GROUP`` = "this optimizer's parameters are not replicated anywhere" (expert shards when the
        # expert-data-parallel size is 1): no reduce-scatter / all-gather at all.  ``None`` = default to the DP group.
        self._shard_group = None if shard_process_group is NO_SHARD_GROUP else shard_process_group
        if shard_process_group is None and dist.is_initialized() and ps.model_parallel_is_initialized() \
                and ps.get_data_parallel_size() > 1:
            self._shard_group = ps.get_data_parallel_group()
        self.sharding_groups = sharding_groups
        self.grad_norm_groups = grad_norm_groups
        self._fused = optimizer_class in (torch.optim.AdamW, torch.optim.Adam) or \
            getattr(optimizer_class, "__name__", "") == "AdamW_FP32OptimParams"
        self._decoupled = optimizer_class is not torch.optim.Adam
        defaults.setdefault("lr", 1e-3)
        if self._fused:
            defaults.setdefault("betas", (0.9, 0.999))
            defaults.setdefault("eps", 1e-8)
            defaults.setdefault("weight_decay", 1e-2 if optimizer_class is torch.optim.AdamW else 0.0)
        super().__init__(params, defaults)
        self._step_count = 0
        self._grad_norm: Optional[torch.Tensor] = None
        self.buckets: List[_Bucket] = []
        self.inner: Optional[torch.optim.Optimizer] = None
        self._scalars: Dict[str, torch.Tensor] = {}
        self._inited = False
        if not lazy_init:
            self.init_zero()

    # ------------------------------------------------------------------------------------------
    def init_zero(self) -> None:
        if self._inited:
            return
        self._inited = True
        for gi, group in enumerate(self.param_groups):
            by_dtype: Dict[Any, List[torch.nn.Parameter]] = {}
            for p in group["params"]:
                if p.requires_grad:
                    by_dtype.setdefault((p.dtype, p.device), []).append(p)
            for (dt, _), plist in by_dtype.items():
                gdt = torch.float32 if (self.use_grad_acc_hook and dt != torch.float32) else None
                b = _Bucket(plist, gi, self._shard_group, self.optimizer_dtype, gdt, symm=self._symm_context(),
                            name=f"zero1:{id(self) & 0xffffff:x}:{len(self.buckets)}")
                if gdt is not None:
                    self._install_grad_acc_hooks(b)
                self.buckets.append(b)
        if not self._fused:
            groups = []
            for gi, group in enumerate(self.param_groups):
                masters = []
                for b in self.buckets:
                    if b.group_idx == gi:
                        b.master_param = torch.nn.Parameter(b.master, requires_grad=True)
                        masters.append(b.master_param)
                g = {k: v for k, v in group.items() if k != "params"}
                g["params"] = masters
                groups.append(g)
            self.inner = self.optimizer_class(groups)
        dev = self.buckets[0].device if self.buckets else torch.device("cpu")
        for name in ("sumsq", "sumsq_dup", "coef", "norm"):
            self._scalars[name] = torch.zeros(1, dtype=torch.float32, device=dev)

    def _symm_context(self):
        """Peer-mapped memory context of the shard group for the fused step, or ``None`` (CPU, single rank, a custom
        shard group, ranks that cannot map each other's memory, ``NXD_FUSED_ZERO1=0``)."""
        import os

        if (os.environ.get("NXD_FUSED_ZERO1", "1") != "1" or not self._fused or self._shard_group is None
                or not torch.cuda.is_available() or not ps.model_parallel_is_initialized()):
            return None
        try:
            if self._shard_group is ps.get_data_parallel_group():
                return ps.get_symm_mem_context("dp")
            if ps.get_expert_model_parallel_size() > 1 and self._shard_group is ps.get_expert_data_parallel_group():
                return ps.get_symm_mem_context("exp_dp")
        except Exception as e:  # noqa: BLE001 - e.g. DP ranks on different nodes: CUDA IPC cannot map them
            logger.warning("fused ZeRO-1 step unavailable (%s); using NCCL collectives", e)
        return None

    def _install_grad_acc_hooks(self, b: _Bucket) -> None:
        """fp32 gradient accumulation (reference ``use_grad_acc_hook`` + ``higher_cc_precision``): params whose
        wgrad GEMM accumulates straight into ``main_grad`` produce no ``.grad``; everything else is folded in
        by a post-accumulate hook."""
        def hook(p):
            if p.grad is not None:
                p.main_grad.add_(p.grad.to(p.main_grad.dtype))
                p.grad = None
        for p in b.params:
            p.register_post_accumulate_grad_hook(hook)

    # ------------------------------------------------------------------------------------------
    @property
    def base_optimizer(self) -> torch.optim.Optimizer:
        """The optimizer over this rank's fp32 master shards, as ``optimizer_class`` (the reference's ZeRO-1 wraps such
        an instance and exposes it under this name).  Non-fused classes really step through it; for the fused AdamW
        path it is a VIEW: its parameters are the master shards and its ``exp_avg / exp_avg_sq`` state tensors alias the
        flat moment buffers, so ``base_optimizer.state_dict()`` and hyper-parameter reads show the live values."""
        self.init_zero()
        if self.inner is not None:
            return self.inner
        view = getattr(self, "_base_view", None)
        if view is None:
            groups = []
            for gi, group in enumerate(self.param_groups):
                g = {k: v for k, v in group.items() if k != "params"}
                g["params"] = []
                for b in self.buckets:
                    if b.group_idx == gi:
                        if getattr(b, "master_param", None) is None:
                            b.master_param = torch.nn.Parameter(b.master, requires_grad=True)
                        g["params"].append(b.master_param)
                groups.append(g)
            view = self.optimizer_class([g for g in groups if g["params"]] or [{"params": []}])
            self._base_view = view
        for b in self.buckets:
            if b.exp_avg is not None and getattr(b, "master_param", None) is not None:
                view.state[b.master_param] = {"step": torch.tensor(float(self._step_count)), "exp_avg": b.exp_avg,
                                              "exp_avg_sq": b.exp_avg_sq}
        return view

    @property
    def grad_norm(self):
        return self._grad_norm

    def zero_grad(self, set_to_none: bool = False) -> None:
        if not self._inited:
            return super().zero_grad(set_to_none=True)
        for b in self.buckets:
            if b.flat_grad.dtype != b.dtype:
                b.flat_grad.zero_()          # fp32 accumulation buffer (main_grad): the GEMM epilogues add into it
                continue
            # No fill kernel: the first gradient of a parameter OVERWRITES its slice of the flat buffer (the wgrad
            # GEMM stores straight into it, other grads are copied in by _sync_grad_views), later ones accumulate.
            for p in b.params:
                p.grad = None
                p._nxd_flat_written = False

    def _norm_groups(self):
        groups = []
        if dist.is_initialized() and ps.model_parallel_is_initialized():
            if ps.get_tensor_model_parallel_size() > 1:
                groups.append(ps.get_tensor_model_parallel_group())
            if ps.get_pipeline_model_parallel_size() > 1:
                groups.append(ps.get_pipeline_model_parallel_group())
        if self._shard_group is not None and _group_size(self._shard_group) > 1:
            groups.append(self._shard_group)
        return groups

    def allreduce_sequence_parallel_gradients(self) -> None:
        """K15 in one collective per bucket: the SP-tagged params are contiguous in the flat grad buffer."""
        if not (dist.is_initialized() and ps.model_parallel_is_initialized()
                and ps.get_tensor_model_parallel_size() > 1):
            return
        g = ps.get_tensor_model_parallel_group()
        self._sync_grad_views()
        for b in self.buckets:
            if b.dup_other_begin > b.sp_begin:
                dist.all_reduce(b.flat_grad[b.sp_begin:b.dup_other_begin], group=g)

    def _sync_grad_views(self) -> None:
        """Re-attach ``p.grad`` to the flat buffer if user code replaced it (``model.zero_grad()`` sets it to
        ``None``; a fresh autograd tensor appears after the next backward)."""
        for b in self.buckets:
            if b.flat_grad.dtype != b.dtype:
                continue
            zero, dst, src = [], [], []
            for p, o in zip(b.params, b.offsets):
                view = b.flat_grad[o:o + p.numel()].view(p.shape)
                g = p.grad
                p._nxd_flat_written = False      # the backward pass that set it is over (see layers._accumulate_wgrad)
                if g is None:
                    zero.append(view)
                elif g.data_ptr() != view.data_ptr():
                    dst.append(view)
                    src.append(g if g.dtype == view.dtype else g.to(view.dtype))
                else:
                    continue
                p.grad = view
            if zero:
                torch._foreach_zero_(zero)
            if dst:
                torch._foreach_copy_(dst, src)

    def _reduce_scatter_grads(self) -> None:
        self._sync_grad_views()
        fused = [b for b in self.buckets if b.symm is not None]
        if fused:
            from ..ops import _lib
            import ctypes as _c

            fused[0].symm.barrier()      # every rank's gradients are complete before anyone pulls them
            # the pull-reduce also squares what it has just summed, so the local part of the gradient norm
            # costs no pass of its own over the fp32 shard (_compute_clip skips these buckets)
            s, sd = self._scalars["sumsq"], self._scalars["sumsq_dup"]
            s.zero_()
            sd.zero_()
            for b in fused:
                dt = 1 if b.flat_grad.dtype == torch.bfloat16 else 0
                if getattr(b, "sumsq_partials", None) is None:
                    nblk = int(_lib.lib().nxd_zero1_reduce_blocks(_c.c_longlong(b.shard_size), _c.c_int(dt)))
                    b.sumsq_partials = torch.zeros(2 * nblk, dtype=torch.float32, device=b.device)
                split_at = min(max(b.sp_begin - b.lo, 0), b.shard_size)
                rc = _lib.lib().nxd_zero1_reduce_shard_sumsq(
                    b.grad_ptrs, _c.c_int(dt), _c.c_longlong(b.lo), _c.c_longlong(b.shard_size), _lib.ptr(b.grad_shard),
                    _c.c_int(b.dp), _c.c_int(b.dp_rank), _c.c_longlong(split_at), _lib.ptr(b.sumsq_partials), _lib.ptr(s),
                    _lib.ptr(sd), _lib.stream_ptr())
                _lib.check(rc, "nxd_zero1_reduce_shard_sumsq")
            self._fused_sumsq = True
        for b in self.buckets:
            if b.dp == 1 or b.symm is not None:
                continue
            src = b.flat_grad
            if self.higher_cc_precision and src.dtype != torch.float32:
                src = src.float()
                if b.grad_shard.dtype != torch.float32:
                    b.grad_shard = torch.zeros(b.shard_size, dtype=torch.float32, device=b.device)
            dist.reduce_scatter_tensor(b.grad_shard, src, op=dist.ReduceOp.SUM, group=b.shard_group)

    def _compute_clip(self) -> Optional[torch.Tensor]:
        """Global grad norm over shard/TP/PP groups; replicated params count 1/tp.  Returns the clip
        coefficient tensor (device) or ``None`` when clipping is off."""
        s, sd = self._scalars["sumsq"], self._scalars["sumsq_dup"]
        fused_done = getattr(self, "_fused_sumsq", False)       # set by _reduce_scatter_grads: s / sd already hold the
        self._fused_sumsq = False                                # sums of the buckets reduced by the fused kernel
        if not fused_done:
            s.zero_()
            sd.zero_()
        for b in self.buckets:
            if fused_done and b.symm is not None:
                continue
            a0, a1 = b.local_range(0, b.sp_begin)
            d0, d1 = b.local_range(b.sp_begin, b.used)
            if a1 > a0:
                optim_ops.sumsq_(b.grad_shard[a0:a1], s, accumulate=True)
            if d1 > d0:
                optim_ops.sumsq_(b.grad_shard[d0:d1], sd, accumulate=True)
        # weights tied across pipeline stages hold the same (summed) gradient on every owning stage: only the first
        # owner counts them (``p.shared`` is set on the others, pipeline/model.py) -- take their part out again
        skip, skip_dup = None, None
        for b in self.buckets:
            for p, o in zip(b.params, b.offsets):
                if not getattr(p, "shared", False):
                    continue
                r0, r1 = b.local_range(o, o + p.numel())
                if r1 <= r0:
                    continue
                if skip is None:
                    skip, skip_dup = torch.zeros_like(s), torch.zeros_like(sd)
                optim_ops.sumsq_(b.grad_shard[r0:r1], skip if o < b.sp_begin else skip_dup, accumulate=True)
        if skip is not None:
            s, sd = s - skip, sd - skip_dup
        tp = ps.get_tensor_model_parallel_size() if ps.model_parallel_is_initialized() else 1
        total = s + sd / float(tp)
        # grads in grad_shard are SUMs over dp (not yet averaged): norm of the averaged grads
        dp = self.buckets[0].dp if self.buckets else 1
        if dp > 1:
            total = total / float(dp * dp)
        for g in self._norm_groups():
            dist.all_reduce(total, group=g)
        self._scalars["total"] = total
        optim_ops.clip_coef_(total, self.max_norm if self.grad_clipping else 0.0, self._scalars["coef"],
                             self._scalars["norm"])
        self._grad_norm = self._scalars["norm"]
        return self._scalars["coef"] if self.grad_clipping else None

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.init_zero()
        self._step_count += 1
        self._reduce_scatter_grads()
        coef = self._compute_clip()
        for b in self.buckets:
            group = self.param_groups[b.group_idx]
            scale = 1.0 / b.dp
            if self._fused:
                b.init_state(self.optimizer_dtype)
                beta1, beta2 = group["betas"]
                out = b.flat_param[b.lo:b.hi]
                if not self._decoupled and group["weight_decay"] != 0.0:
                    b.grad_shard.add_(b.master.to(b.grad_shard.dtype), alpha=group["weight_decay"] / scale)
                if b.symm is not None:
                    # AdamW on the shard; the bf16 parameters are stored into every rank's flat buffer (= all-gather)
                    from ..ops import _lib
                    import ctypes as _c

                    rc = _lib.lib().nxd_zero1_adamw_allgather(
                        _lib.ptr(b.master), _lib.ptr(b.exp_avg), _lib.ptr(b.exp_avg_sq), _lib.ptr(b.grad_shard),
                        b.param_ptrs, _c.c_longlong(b.lo), _c.c_longlong(b.shard_size), _c.c_float(group["lr"]),
                        _c.c_float(beta1), _c.c_float(beta2), _c.c_float(group["eps"]),
                        _c.c_float(group["weight_decay"] if self._decoupled else 0.0), _c.c_int(self._step_count),
                        _c.c_float(scale), _lib.ptr(coef), _c.c_int(b.dp), _c.c_int(b.dp_rank), _lib.stream_ptr())
                    _lib.check(rc, "nxd_zero1_adamw_allgather")
                    continue
                optim_ops.adamw_step_(b.master, b.exp_avg, b.exp_avg_sq, b.grad_shard,
                                      None if out.dtype == b.master.dtype else out, group["lr"], beta1, beta2,
                                      group["eps"], group["weight_decay"] if self._decoupled else 0.0,
                                      self._step_count, grad_scale=scale, clip_coef=coef)
                if out.dtype == b.master.dtype:
                    out.copy_(b.master)
            else:
                g = b.grad_shard.to(b.master.dtype) * scale
                if coef is not None:
                    g = g * coef
                b.master_param.grad = g
        if not self._fused:
            for ig, og in zip(self.inner.param_groups, self.param_groups):
                for k, v in og.items():
                    if k != "params":
                        ig[k] = v
            self.inner.step()
            for b in self.buckets:
                b.flat_param[b.lo:b.hi].copy_(b.master_param.data)
                b.master_param.grad = None
        pushed = None
        for b in self.buckets:
            if b.dp > 1 and b.symm is not None and self._fused:
                pushed = b.symm
            elif b.dp > 1:
                dist.all_gather_into_tensor(b.flat_param, b.flat_param[b.lo:b.hi].clone(), group=b.shard_group)
        if pushed is not None:
            pushed.barrier()             # every rank's parameter slices have landed before the next forward
        return loss

    # ------------------------------------------------------------------------------------------
    # checkpoint state (reference keys; flat-slice layout)
    # ------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        self.init_zero()
        state, base_state, shape_info = {}, {}, {}
        idx = 0
        layout = []
        for bi, b in enumerate(self.buckets):
            for p, o in zip(b.params, b.offsets):
                a0, a1 = b.local_range(o, o + p.numel())
                shape_info[idx] = tuple(p.shape)
                base_state[idx] = b.master[a0:a1].clone()
                if self._fused and b.exp_avg is not None:
                    state[idx] = {"step": torch.tensor(float(self._step_count)),
                                  "exp_avg": b.exp_avg[a0:a1].clone(), "exp_avg_sq": b.exp_avg_sq[a0:a1].clone()}
                layout.append({"bucket": bi, "offset": o, "numel": p.numel(),
                               "begin": max(o, b.lo) - o if a1 > a0 else 0, "len": a1 - a0})
                idx += 1
        sd = {
            "state": state, "base_state": base_state, "shape_info": shape_info,
            "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
            "layout": {"kind": "flat_contiguous", "params": layout, "dp_size": self.buckets[0].dp if self.buckets else 1,
                       "dp_rank": self.buckets[0].dp_rank if self.buckets else 0, "step": self._step_count},
        }
        if not self._fused and self.inner is not None:
            sd["inner"] = self.inner.state_dict()
        return sd

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.init_zero()
        self._step_count = int(sd.get("layout", {}).get("step", 0))
        idx = 0
        for b in self.buckets:
            if self._fused:
                b.init_state(self.optimizer_dtype)
            for p, o in zip(b.params, b.offsets):
                a0, a1 = b.local_range(o, o + p.numel())
                if a1 > a0:
                    b.master[a0:a1].copy_(sd["base_state"][idx].to(b.master.device))
                    st = sd["state"].get(idx)
                    if st is not None and self._fused:
                        b.exp_avg[a0:a1].copy_(st["exp_avg"].to(b.master.device))
                        b.exp_avg_sq[a0:a1].copy_(st["exp_avg_sq"].to(b.master.device))
                        self._step_count = max(self._step_count, int(float(st["step"])))
                idx += 1
        for g, saved in zip(self.param_groups, sd["param_groups"]):
            g.update({k: v for k, v in saved.items() if k != "params"})
        if not self._fused and "inner" in sd:
            self.inner.load_state_dict(sd["inner"])

    # deprecated sharded save/load of the reference (``:102-155``), kept for API parity
    def save_sharded_state_dict(self, output_dir: str, num_workers_per_step: int = 8) -> None:
        import os

        os.makedirs(os.path.join(output_dir, "optim"), exist_ok=True)
        dp_rank = ps.get_data_parallel_rank() if ps.model_parallel_is_initialized() else 0
        tp_rank = ps.get_tensor_model_parallel_rank() if ps.model_parallel_is_initialized() else 0
        torch.save(self.state_dict(), os.path.join(output_dir, "optim", f"optim.dp_rank_{dp_rank:02d}.tp_rank_{tp_rank:02d}"))

    def load_sharded_state_dict(self, output_dir: str, num_workers_per_step: int = 8) -> None:
        import os

        dp_rank = ps.get_data_parallel_rank() if ps.model_parallel_is_initialized() else 0
        tp_rank = ps.get_tensor_model_parallel_rank() if ps.model_parallel_is_initialized() else 0
        path = os.path.join(output_dir, "optim", f"optim.dp_rank_{dp_rank:02d}.tp_rank_{tp_rank:02d}")
        self.load_state_dict(torch.load(path, map_location="cpu", weights_only=False))


class NeuronEPZero1Optimizer(torch.optim.Optimizer):
    """Expert-parallel aware ZeRO-1 (reference ``zero_redundancy_optimizer.py:158-362``): non-expert params are
    sharded over the full DP group, expert params (``expert_model_parallel=True``) over the expert-DP group;
    one global grad norm over both."""

    def __init__(self, params, optimizer_class=torch.optim.AdamW, **kw):
        params = list(params)
        if params and isinstance(params[0], dict):
            groups = params
        else:
            groups = [{"params": params}]
        ep_groups, nonep_groups = [], []
        for g in groups:
            ep = [p for p in g["params"] if getattr(p, "expert_model_parallel", False)]
            ne = [p for p in g["params"] if not getattr(p, "expert_model_parallel", False)]
            rest = {k: v for k, v in g.items() if k != "params"}
            if ep:
                ep_groups.append(dict(rest, params=ep))
            if ne:
                nonep_groups.append(dict(rest, params=ne))
        self.grad_clipping = kw.get("grad_clipping", True)
        self.max_norm = kw.get("max_norm", 1.0) or 1.0
        inner_kw = dict(kw, grad_clipping=False)
        # expert shards differ between EP ranks: they may only ever be reduced / gathered inside the expert-DP group,
        # and not at all when that group has one member (falling back to the DP group would mix different experts)
        exp_dp = ps.get_expert_data_parallel_group() if (
            dist.is_initialized() and ps.model_parallel_is_initialized()
            and ps.get_expert_data_parallel_size() > 1) else NO_SHARD_GROUP
        self.non_ep_zero_optimizer = NeuronZero1Optimizer(nonep_groups, optimizer_class, **inner_kw) \
            if nonep_groups else None
        self.ep_zero_optimizer = NeuronZero1Optimizer(ep_groups, optimizer_class, shard_process_group=exp_dp,
                                                      **inner_kw) if ep_groups else None
        self._ep_size = ps.get_expert_model_parallel_size() if ps.model_parallel_is_initialized() else 1
        self._opts = [o for o in (self.non_ep_zero_optimizer, self.ep_zero_optimizer) if o is not None]
        defaults = {k: v for k, v in kw.items() if k in ("lr", "betas", "eps", "weight_decay")}
        defaults.setdefault("lr", 1e-3)
        torch.optim.Optimizer.__init__(self, groups, defaults)
        # expose the inner groups so LR schedulers drive both
        self.param_groups = [g for o in self._opts for g in o.param_groups]
        self._grad_norm = None

    @property
    def sharding_groups(self):
        """Rank meshes the two inner optimizers shard over: ``(data-parallel, expert-data-parallel)`` (reference
        ``:204-214``)."""
        if not (dist.is_initialized() and ps.model_parallel_is_initialized()):
            return None
        return ps.get_data_parallel_group(as_list=True), ps.get_expert_data_parallel_group(as_list=True)

    @sharding_groups.setter
    def sharding_groups(self, value) -> None:
        """Accepted for compatibility with the base class, which assigns it; the groups come from ``parallel_state``."""

    @property
    def grad_norm(self):
        return self._grad_norm

    def zero_grad(self, set_to_none: bool = False) -> None:
        for o in self._opts:
            o.zero_grad(set_to_none)

    def allreduce_sequence_parallel_gradients(self) -> None:
        for o in self._opts:
            o.allreduce_sequence_parallel_gradients()

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        # expert grads: the EP ranks hold different experts, so the expert-DP reduce-scatter averages over
        # exp_dp only while the batch was split over dp = exp_dp * ep -> scale by 1/ep (reference :231-268)
        if self.ep_zero_optimizer is not None and self._ep_size > 1:
            self.ep_zero_optimizer._sync_grad_views()
            for b in self.ep_zero_optimizer.buckets:
                b.flat_grad.mul_(1.0 / self._ep_size)
        for o in self._opts:
            o._step_count += 1
            o._reduce_scatter_grads()
        total = None
        for o in self._opts:
            o._compute_clip()
            t = o._scalars["total"].clone()
            if o is self.ep_zero_optimizer and self._ep_size > 1:
                dist.all_reduce(t, group=ps.get_expert_model_parallel_group())
            total = t if total is None else total + t
        coef = None
        if total is not None:
            norm = total.sqrt()
            self._grad_norm = norm
            if self.grad_clipping:
                coef = torch.clamp(self.max_norm / (norm + 1e-6), max=1.0)
        for o in self._opts:
            o._step_count -= 1
            o._external_coef = coef
            _step_with_coef(o, coef)
        return loss

    def state_dict(self):
        return {"non_ep": self.non_ep_zero_optimizer.state_dict() if self.non_ep_zero_optimizer else None,
                "ep": self.ep_zero_optimizer.state_dict() if self.ep_zero_optimizer else None}

    def load_state_dict(self, sd):
        if self.non_ep_zero_optimizer is not None and sd.get("non_ep") is not None:
            self.non_ep_zero_optimizer.load_state_dict(sd["non_ep"])
        if self.ep_zero_optimizer is not None and sd.get("ep") is not None:
            self.ep_zero_optimizer.load_state_dict(sd["ep"])


def _step_with_coef(opt: NeuronZero1Optimizer, coef: Optional[torch.Tensor]) -> None:
    """Run the update phase of ``opt.step()`` with an externally computed clip coefficient (grads already
    reduce-scattered)."""
    saved_rs, saved_clip = opt._reduce_scatter_grads, opt._compute_clip
    opt._reduce_scatter_grads = lambda: None
    opt._compute_clip = lambda: coef
    try:
        opt.step()
    finally:
        opt._reduce_scatter_grads, opt._compute_clip = saved_rs, saved_clip
