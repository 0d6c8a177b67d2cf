# Sourced by the launch scripts: builds the torchrun argument list for one node or a SLURM allocation.
# B200 nodes have 8 GPUs; one process per GPU.  Rendezvous on 127.0.0.1 for a single node (container host names may not
# resolve), on the first SLURM host otherwise.
: "${GPUS_PER_NODE:=8}"
REPO_ROOT=$(cd -- "$(dirname -- "${BASH_SOURCE[0]}")/../../.." &> /dev/null && pwd)
EXAMPLES=$REPO_ROOT/examples/training
CONFIGS=$EXAMPLES/configs
NNODES=1
NODE_RANK=0
MASTER_ADDR=127.0.0.1
: "${MASTER_PORT:=29500}"
if [ -n "${SLURM_NTASKS:-}" ]; then
    NNODES=$SLURM_NTASKS
    NODE_RANK=$SLURM_NODEID
    MASTER_ADDR=$(scontrol show hostnames "$SLURM_JOB_NODELIST" | head -n 1)
fi
WORLD_GPUS=$((GPUS_PER_NODE * NNODES))
DISTRIBUTED_ARGS="--nproc-per-node $GPUS_PER_NODE --nnodes $NNODES --node-rank $NODE_RANK --master-addr $MASTER_ADDR --master-port $MASTER_PORT"
# NCCL over NVLink/NVSwitch inside the node; leave NVLS on (in-switch reductions) unless debugging
export NCCL_DEBUG=${NCCL_DEBUG:-WARN}
export TORCH_NCCL_ASYNC_ERROR_HANDLING=1
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-8}
# PERF_TEST=1: 100 steps, no checkpoints (the reference's scripts use the same switch)
if [ "${PERF_TEST:-0}" -gt 0 ]; then STEPS_OVERRIDE=100; CKPT_FREQ_OVERRIDE=0; fi
LOG_DIR=${LOG_DIR:-logs}
mkdir -p "$LOG_DIR"
