"""Topology and process groups.

Parity target: reference ``src/neuronx_distributed/parallel_layers/parallel_state.py:60-766``.

The world is a row-major rank grid ``[PP, DP, TP]`` (non-expert view) which is the
same memory viewed as ``[PP, DP_exp, EP, TP]`` inside expert regions; TP is the
stride-1 dimension so a TP group lands on one NVSwitch domain (8 x B200).

B200-first differences from the reference:
  * groups are real ``torch.distributed`` groups on NCCL (GPU) or gloo (CPU tests);
    the XLA "replica group list" twin is kept only as plain python lists
    (``as_list=True``) for API compatibility and for the checkpoint code;
  * every group may carry a *symmetric-memory context* (peer-mapped buffers +
    signal pads, see ``ops/symm_mem.py``) that the fused AG->GEMM / GEMM->RS /
    ZeRO-1 kernels use instead of NCCL; it is attached lazily by
    ``get_symm_mem_context``;
  * PP send/recv uses true NCCL p2p, so the parity-paired 2-rank send/recv groups
    of the reference (``parallel_state.py:368-385``) are kept only as rank lists.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from ..utils.logger import get_logger

logger = get_logger()


class _GroupInfo:
    """One family of groups (e.g. all TP groups): my group, my rank list, full mesh."""

    __slots__ = ("group", "ranks", "mesh")

    def __init__(self, group, ranks, mesh):
        self.group = group
        self.ranks = ranks
        self.mesh = mesh


_TP: Optional[_GroupInfo] = None
_PP: Optional[_GroupInfo] = None
_DP: Optional[_GroupInfo] = None
_EP: Optional[_GroupInfo] = None
_EXP_DP: Optional[_GroupInfo] = None
_PP_GLOO = None
_NEXT_RANK_LIST: Optional[List[List[int]]] = None
_PREV_RANK_LIST: Optional[List[List[int]]] = None
_KV_SHARED: Dict[int, _GroupInfo] = {}
_CUSTOM_PGS: Dict[tuple, Any] = {}
_SYMM_CTX: Dict[str, Any] = {}

# Manual overrides used to mock a topology inside one process
# (reference: ``_MPU_*`` variables, parallel_state.py:477-538).
_MPU_TENSOR_MODEL_PARALLEL_WORLD_SIZE: Optional[int] = None
_MPU_TENSOR_MODEL_PARALLEL_RANK: Optional[int] = None
_MPU_EXPERT_MODEL_PARALLEL_WORLD_SIZE: Optional[int] = None
_MPU_EXPERT_MODEL_PARALLEL_RANK: Optional[int] = None
_MPU_PIPELINE_MODEL_PARALLEL_WORLD_SIZE: Optional[int] = None
_MPU_PIPELINE_MODEL_PARALLEL_RANK: Optional[int] = None
_MPU_DATA_PARALLEL_WORLD_SIZE: Optional[int] = None
_MPU_DATA_PARALLEL_RANK: Optional[int] = None


def compute_rank_meshes(world_size: int, tp: int, pp: int, ep: int) -> Dict[str, List[List[int]]]:
    """Pure function: all group rank lists for a (world, tp, pp, ep) topology.

    Mirrors the group maths of reference parallel_state.py:246-360 (TP contiguous,
    DP stride TP, PP stride DP*TP; expert view splits DP into DP_exp x EP).
    The Trn1-only TP=4 interleaved layout (``:218-244``) is a NeuronLink wiring
    artefact and is intentionally not reproduced: on NVSwitch every peer is
    equidistant.
    """
    if world_size % (tp * pp) != 0:
        raise RuntimeError(
            f"`world_size` ({world_size}) is not divisible by tensor_model_parallel_size ({tp}) x "
            f"pipeline_model_parallel_size ({pp})"
        )
    dp = world_size // (tp * pp)
    if world_size % (tp * pp * ep) != 0:
        raise RuntimeError(
            f"invalid implied expert data parallel degree: `world_size` ({world_size}) is not divisible by "
            f"tp ({tp}) x pp ({pp}) x ep ({ep})"
        )
    dp_exp = world_size // (tp * pp * ep)
    ranks = torch.arange(world_size)
    nonexp = ranks.reshape(pp, dp, tp)
    exp = ranks.reshape(pp, dp_exp, ep, tp)
    meshes = {
        "tp": nonexp.reshape(-1, tp).tolist(),
        "dp": nonexp.permute(0, 2, 1).reshape(-1, dp).tolist(),
        "pp": nonexp.permute(1, 2, 0).reshape(-1, pp).tolist(),
        "ep": exp.permute(0, 1, 3, 2).reshape(-1, ep).tolist(),
        "exp_dp": exp.permute(0, 2, 3, 1).reshape(-1, dp_exp).tolist(),
    }
    return meshes


def _build_groups(mesh: List[List[int]], rank: int, backend: Optional[str]) -> _GroupInfo:
    """Create one ``ProcessGroup`` per row of ``mesh`` (all ranks must call this
    in the same order) and remember the one ``rank`` belongs to
    (reference ``_build_and_assign_groups`` parallel_state.py:388-422)."""
    mine, mine_ranks = None, None
    for row in mesh:
        pg = dist.new_group(ranks=row, backend=backend)
        if rank in row:
            mine, mine_ranks = pg, row
    assert mine_ranks is not None
    return _GroupInfo(mine, mine_ranks, mesh)


def initialize_model_parallel(
    tensor_model_parallel_size: int = 1,
    pipeline_model_parallel_size: int = 1,
    expert_model_parallel_size: int = 1,
) -> None:
    """Build TP / PP / DP / EP / EXP-DP groups (reference parallel_state.py:60-385)."""
    global _TP, _PP, _DP, _EP, _EXP_DP, _NEXT_RANK_LIST, _PREV_RANK_LIST
    assert dist.is_initialized(), "torch.distributed must be initialised first"
    if model_parallel_is_initialized():
        raise RuntimeError("model parallel groups are already initialized")
    world_size = dist.get_world_size()
    rank = dist.get_rank()
    tp = min(tensor_model_parallel_size, world_size)
    pp = min(pipeline_model_parallel_size, world_size)
    ep = min(expert_model_parallel_size, world_size)
    meshes = compute_rank_meshes(world_size, tp, pp, ep)
    if rank == 0:
        logger.info(
            "initialize_model_parallel: world=%d tp=%d pp=%d dp=%d ep=%d exp_dp=%d",
            world_size, tp, pp, world_size // (tp * pp), ep, world_size // (tp * pp * ep),
        )

    # Communicator warm-up (reference does a dummy world all-reduce, :276-280).
    if os.environ.get("NXD_SKIP_WARMUP_ALLREDUCE", "0") != "1":
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu"
        t = torch.ones(1, device=dev)
        dist.all_reduce(t)

    backend = None  # inherit default backend (nccl on GPU, gloo on CPU)
    _DP = _build_groups(meshes["dp"], rank, backend)
    _EXP_DP = _build_groups(meshes["exp_dp"], rank, backend)
    _TP = _build_groups(meshes["tp"], rank, backend)
    _EP = _build_groups(meshes["ep"], rank, backend)
    _PP = _build_groups(meshes["pp"], rank, backend)

    # next/prev neighbour lists (rank pairs) for PP p2p (reference :368-385)
    nxt, prv = [], []
    for row in meshes["pp"]:
        for i in range(len(row)):
            nxt.append([row[i], row[(i + 1) % len(row)]])
            prv.append([row[i], row[(i - 1) % len(row)]])
    _NEXT_RANK_LIST, _PREV_RANK_LIST = nxt, prv


def model_parallel_is_initialized() -> bool:
    return _TP is not None and _PP is not None and _DP is not None


class _Solo(_GroupInfo):
    """Stand-in used when neither torch.distributed nor model parallel is up:
    a single-process run behaves as tp=pp=dp=ep=1 (handy for 1-GPU runs/tests)."""

    def __init__(self):
        super().__init__(None, [0], [[0]])


def _need(info: Optional[_GroupInfo], what: str) -> _GroupInfo:
    if info is None and not dist.is_initialized():
        return _Solo()
    assert info is not None, f"{what} group is not initialized"
    return info


def _my_rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def get_tensor_model_parallel_group(as_list: bool = False):
    g = _need(_TP, "tensor model parallel")
    return g.mesh if as_list else g.group


def get_expert_model_parallel_group(as_list: bool = False):
    g = _need(_EP, "expert model parallel")
    return g.mesh if as_list else g.group


def get_data_parallel_group(as_list: bool = False):
    g = _need(_DP, "data parallel")
    return g.mesh if as_list else g.group


def get_expert_data_parallel_group(as_list: bool = False):
    g = _need(_EXP_DP, "expert data parallel")
    return g.mesh if as_list else g.group


def get_pipeline_model_parallel_group(as_list: bool = False):
    g = _need(_PP, "pipeline model parallel")
    return g.mesh if as_list else g.group


def get_tensor_model_parallel_ranks() -> List[int]:
    return list(_need(_TP, "tensor model parallel").ranks)


def get_data_parallel_ranks() -> List[int]:
    return list(_need(_DP, "data parallel").ranks)


def get_pipeline_model_parallel_ranks() -> List[int]:
    return list(_need(_PP, "pipeline model parallel").ranks)


def get_expert_model_parallel_ranks() -> List[int]:
    return list(_need(_EP, "expert model parallel").ranks)


# ---- sizes / ranks (with manual overrides for mocked topologies) -------------

def get_tensor_model_parallel_size() -> int:
    if _MPU_TENSOR_MODEL_PARALLEL_WORLD_SIZE is not None:
        return _MPU_TENSOR_MODEL_PARALLEL_WORLD_SIZE
    return len(_need(_TP, "tensor model parallel").ranks)


def set_tensor_model_parallel_size(world_size: Optional[int]) -> None:
    global _MPU_TENSOR_MODEL_PARALLEL_WORLD_SIZE
    _MPU_TENSOR_MODEL_PARALLEL_WORLD_SIZE = world_size


def set_tensor_model_parallel_rank(rank: Optional[int]) -> None:
    global _MPU_TENSOR_MODEL_PARALLEL_RANK
    _MPU_TENSOR_MODEL_PARALLEL_RANK = rank


def get_tensor_model_parallel_rank() -> int:
    if _MPU_TENSOR_MODEL_PARALLEL_RANK is not None:
        return _MPU_TENSOR_MODEL_PARALLEL_RANK
    return _need(_TP, "tensor model parallel").ranks.index(_my_rank())


def get_tensor_model_parallel_src_rank() -> int:
    """Global rank of the first rank in my TP group."""
    if _TP is None:
        return 0
    return _TP.ranks[0]


def set_expert_model_parallel_size(world_size: Optional[int]) -> None:
    global _MPU_EXPERT_MODEL_PARALLEL_WORLD_SIZE
    _MPU_EXPERT_MODEL_PARALLEL_WORLD_SIZE = world_size


def get_expert_model_parallel_size() -> int:
    if _MPU_EXPERT_MODEL_PARALLEL_WORLD_SIZE is not None:
        return _MPU_EXPERT_MODEL_PARALLEL_WORLD_SIZE
    return len(_need(_EP, "expert model parallel").ranks)


def set_expert_model_parallel_rank(rank: Optional[int]) -> None:
    global _MPU_EXPERT_MODEL_PARALLEL_RANK
    _MPU_EXPERT_MODEL_PARALLEL_RANK = rank


def get_expert_model_parallel_rank() -> int:
    if _MPU_EXPERT_MODEL_PARALLEL_RANK is not None:
        return _MPU_EXPERT_MODEL_PARALLEL_RANK
    return _need(_EP, "expert model parallel").ranks.index(_my_rank())


def get_data_parallel_src_rank() -> int:
    return _need(_DP, "data parallel").ranks[0]


def set_data_parallel_size(world_size: Optional[int]) -> None:
    global _MPU_DATA_PARALLEL_WORLD_SIZE
    _MPU_DATA_PARALLEL_WORLD_SIZE = world_size


def set_data_parallel_rank(rank: Optional[int]) -> None:
    global _MPU_DATA_PARALLEL_RANK
    _MPU_DATA_PARALLEL_RANK = rank


def get_data_parallel_size() -> int:
    if _MPU_DATA_PARALLEL_WORLD_SIZE is not None:
        return _MPU_DATA_PARALLEL_WORLD_SIZE
    return len(_need(_DP, "data parallel").ranks)


def get_data_parallel_rank() -> int:
    if _MPU_DATA_PARALLEL_RANK is not None:
        return _MPU_DATA_PARALLEL_RANK
    return _need(_DP, "data parallel").ranks.index(_my_rank())


def get_expert_data_parallel_size() -> int:
    if _MPU_DATA_PARALLEL_WORLD_SIZE is not None and _EXP_DP is None:
        return _MPU_DATA_PARALLEL_WORLD_SIZE // get_expert_model_parallel_size()
    return len(_need(_EXP_DP, "expert data parallel").ranks)


def get_expert_data_parallel_rank() -> int:
    if _MPU_DATA_PARALLEL_RANK is not None and _EXP_DP is None:
        return _MPU_DATA_PARALLEL_RANK // get_expert_model_parallel_size()
    return _need(_EXP_DP, "expert data parallel").ranks.index(_my_rank())


def set_pipeline_model_parallel_size(world_size: Optional[int]) -> None:
    global _MPU_PIPELINE_MODEL_PARALLEL_WORLD_SIZE
    _MPU_PIPELINE_MODEL_PARALLEL_WORLD_SIZE = world_size


def set_pipeline_model_parallel_rank(rank: Optional[int]) -> None:
    global _MPU_PIPELINE_MODEL_PARALLEL_RANK
    _MPU_PIPELINE_MODEL_PARALLEL_RANK = rank


def get_pipeline_model_parallel_rank() -> int:
    if _MPU_PIPELINE_MODEL_PARALLEL_RANK is not None:
        return _MPU_PIPELINE_MODEL_PARALLEL_RANK
    return _need(_PP, "pipeline model parallel").ranks.index(_my_rank())


def get_pipeline_model_parallel_size() -> int:
    if _MPU_PIPELINE_MODEL_PARALLEL_WORLD_SIZE is not None:
        return _MPU_PIPELINE_MODEL_PARALLEL_WORLD_SIZE
    return len(_need(_PP, "pipeline model parallel").ranks)


def get_pipeline_model_parallel_sr_group(parity: bool):
    """The reference pairs PP ranks into 2-rank groups by parity because its
    send/recv is an all-gather (parallel_state.py:580-591).  With NCCL p2p the
    whole PP group is the communicator, so both parities map to it."""
    return get_pipeline_model_parallel_group()


def get_next_rank_group(as_list: bool = False):
    assert _NEXT_RANK_LIST is not None
    return _NEXT_RANK_LIST if as_list else get_pipeline_model_parallel_group()


def get_prev_rank_group(as_list: bool = False):
    assert _PREV_RANK_LIST is not None
    return _PREV_RANK_LIST if as_list else get_pipeline_model_parallel_group()


def get_pipeline_model_parallel_next_rank() -> int:
    ranks = _need(_PP, "pipeline model parallel").ranks
    return ranks[(get_pipeline_model_parallel_rank() + 1) % len(ranks)]


def get_pipeline_model_parallel_prev_rank() -> int:
    ranks = _need(_PP, "pipeline model parallel").ranks
    return ranks[(get_pipeline_model_parallel_rank() - 1) % len(ranks)]


def is_pipeline_first_stage() -> bool:
    return get_pipeline_model_parallel_rank() == 0


def is_pipeline_last_stage() -> bool:
    return get_pipeline_model_parallel_rank() == get_pipeline_model_parallel_size() - 1


# ---- KV-shared groups (GQA replication, reference qkv_linear.py:34-72) -------

def get_kv_shared_group(kv_shared_group_size: int, as_list: bool = False):
    """Ranks inside one TP group that hold a replica of the same KV head.

    With ``kv_size_multiplier = m`` and TP size ``t`` the KV weight is tiled
    ``m`` times along the head axis, so TP rank ``r`` holds KV head
    ``r % (t/m)``; ranks ``{j, j + t/m, j + 2t/m, ...}`` share a head.
    """
    m = kv_shared_group_size
    if m in _KV_SHARED:
        g = _KV_SHARED[m]
        return g.mesh if as_list else g.group
    tp_mesh = _need(_TP, "tensor model parallel").mesh
    tp = len(tp_mesh[0])
    assert tp % m == 0, f"tp size {tp} not divisible by kv_size_multiplier {m}"
    stride = tp // m
    mesh = []
    for row in tp_mesh:
        for j in range(stride):
            mesh.append([row[j + k * stride] for k in range(m)])
    info = _build_groups(mesh, dist.get_rank(), None)
    _KV_SHARED[m] = info
    return info.mesh if as_list else info.group


def get_kv_shared_group_size_rank(kv_shared_group_size: int) -> int:
    get_kv_shared_group(kv_shared_group_size)
    return _KV_SHARED[kv_shared_group_size].ranks.index(_my_rank())


# ---- misc --------------------------------------------------------------------

def destroy_model_parallel() -> None:
    """Drop all groups and overrides (reference parallel_state.py:625-652)."""
    global _TP, _PP, _DP, _EP, _EXP_DP, _PP_GLOO, _NEXT_RANK_LIST, _PREV_RANK_LIST
    for ctx in list(_SYMM_CTX.values()):
        try:
            ctx.close()
        except Exception:  # pragma: no cover - teardown best effort
            pass
    _SYMM_CTX.clear()
    _TP = _PP = _DP = _EP = _EXP_DP = None
    _PP_GLOO = None
    _NEXT_RANK_LIST = _PREV_RANK_LIST = None
    _KV_SHARED.clear()
    _CUSTOM_PGS.clear()
    try:
        from ..pipeline import comm as _pp_comm

        _pp_comm.reset_p2p_groups()
    except Exception:  # pragma: no cover
        pass
    for setter in (
        set_tensor_model_parallel_size, set_tensor_model_parallel_rank,
        set_expert_model_parallel_size, set_expert_model_parallel_rank,
        set_pipeline_model_parallel_size, set_pipeline_model_parallel_rank,
        set_data_parallel_size, set_data_parallel_rank,
    ):
        setter(None)


def is_tcp_store_available() -> bool:
    return dist.is_initialized() and dist.distributed_c10d._get_default_store() is not None


def get_tcp_store():
    assert is_tcp_store_available()
    return dist.distributed_c10d._get_default_store()


def initialize_pp_gloo_groups() -> None:
    """CPU side-channel group per PP group (reference :667-676)."""
    global _PP_GLOO
    if _PP_GLOO is not None:
        return
    mine = None
    for row in _need(_PP, "pipeline model parallel").mesh:
        pg = dist.new_group(ranks=row, backend="gloo")
        if dist.get_rank() in row:
            mine = pg
    _PP_GLOO = mine


def get_pp_gloo_group():
    assert _PP_GLOO is not None, "pp gloo groups are not initialized"
    return _PP_GLOO


def is_global_rank_zero() -> bool:
    return (not dist.is_initialized()) or dist.get_rank() == 0


def create_pg_with_ranks(ranks: List[int]):
    """Group over an arbitrary rank list, replicated across PP-equivalent peers
    (reference :693-725: every rank must participate in every new_group call,
    so the list is all-gathered and de-duplicated first)."""
    world = dist.get_world_size()
    gathered: List[Any] = [None] * world
    dist.all_gather_object(gathered, list(ranks))
    uniq = sorted({tuple(r) for r in gathered})
    mine = None
    for r in uniq:
        key = tuple(r)
        if key not in _CUSTOM_PGS:
            _CUSTOM_PGS[key] = dist.new_group(ranks=list(r))
        if tuple(ranks) == key:
            mine = _CUSTOM_PGS[key]
    return mine


def gather_python_object(obj: Any, group) -> List[Any]:
    out: List[Any] = [None] * dist.get_world_size(group=group)
    dist.all_gather_object(out, obj, group=group)
    return out


def rmsg(msg: str) -> str:
    """Prefix ``msg`` with ``[rank_G_ppP_tpT_dpD]`` (reference :740-758)."""
    try:
        return (
            f"[rank_{dist.get_rank()}_pp{get_pipeline_model_parallel_rank()}"
            f"_tp{get_tensor_model_parallel_rank()}_dp{get_data_parallel_rank()}] {msg}"
        )
    except Exception:
        return f"[rank_unknown] {msg}"


def rmsg_ep(msg: str) -> str:
    try:
        return (
            f"[rank_{dist.get_rank()}_pp{get_pipeline_model_parallel_rank()}"
            f"_tp{get_tensor_model_parallel_rank()}_dp{get_data_parallel_rank()}"
            f"_ep{get_expert_model_parallel_rank()}_expdp{get_expert_data_parallel_rank()}] {msg}"
        )
    except Exception:
        return f"[rank_unknown] {msg}"


# ---- symmetric-memory contexts (B200 addition) ---------------

def get_symm_mem_context(kind: str = "tp"):
    """Peer-mapped buffer context for a group family ("tp", "dp", "ep").

    Built lazily the first time a fused kernel needs it; returns ``None`` on CPU
    or when the group has a single rank.  This replaces the reference's
    ``try_set_nki_parallel_state`` (parallel_state.py:425-444), which injects TP
    info into NKI kernels: here the context carries peer pointers, rank and world
    for the sm_100a kernels.
    """
    if kind in _SYMM_CTX:
        return _SYMM_CTX[kind]
    info = {"tp": _TP, "dp": _DP, "ep": _EP, "exp_dp": _EXP_DP}[kind]
    if info is None or len(info.ranks) == 1 or not torch.cuda.is_available():
        return None
    from ..ops.symm_mem import SymmMemContext

    ctx = SymmMemContext(info.group, info.ranks)
    _SYMM_CTX[kind] = ctx
    return ctx


def try_set_nki_parallel_state(parallel_state=None) -> bool:
    """The reference hands the TP topology to its NKI kernels (``parallel_state.py:425-444``).  The sm_100a kernels get
    theirs through the symmetric-memory context (peer pointers, rank, world); this call builds it eagerly and reports
    whether the fused kernels are available."""
    try:
        return get_symm_mem_context("tp") is not None
    except Exception:  # noqa: BLE001 - ranks that cannot map each other's memory
        return False
