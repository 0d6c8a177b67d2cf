#!/bin/bash
# Everything that was added without GPU time, in the order it should first meet hardware.  Each step runs under its own
# timeout and writes a log to gpurun_out/; a failing step does not stop the later ones.  Typical use:
#   gpurun --gpus 2 --timeout 1500 -- 'bash tools/gpu_checklist.sh'
# Steps that need 2 GPUs are skipped on a 1-GPU box.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
NGPU=$(python -c 'import torch; print(torch.cuda.device_count())' 2>/dev/null || echo 0)
echo "GPUs visible: $NGPU"
SUMMARY="$OUT/gpu_checklist_summary.txt"
: > "$SUMMARY"

step() {   # name timeout command...
    local name=$1 limit=$2; shift 2
    echo "== $name"
    timeout "$limit" "$@" > "$OUT/$name.log" 2>&1
    local rc=$?
    echo "$name rc=$rc  $(tail -n 1 "$OUT/$name.log" | cut -c1-160)" | tee -a "$SUMMARY"
}

# 1. opt-in numerics tests of kernels that have never run
NXD_UNVERIFIED=1 step i8_skinny_numerics 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k skinny_gemm_i8 -p no:cacheprovider
# 2. inference features written on CPU: speculation verify graph, page pool + continuous batching, Medusa, int8 runner
step spec_example 300 python examples/inference/run_llama_speculative.py --model llama3.2-1b --draft_model llama3.2-1b \
    --draft_layers 4 --max_context_length 512 --seq_len 1024 --max_new_tokens 64 --speculation_length 4
step continuous_batching 300 python examples/inference/run_continuous_batching.py --model llama3.2-1b --max_context_length 512 \
    --seq_len 2048 --max_new_tokens 64 --max_running 8 --requests 32 --no_cuda_graphs
step quantized_runner 300 python examples/inference/run_llama_quantized.py --model llama3.2-1b --max_context_length 512 \
    --seq_len 1024 --max_new_tokens 32 --benchmark
NXD_INT8_SKINNY=1 step quantized_runner_i8_kernel 300 python examples/inference/run_llama_quantized.py --model llama3.2-1b \
    --max_context_length 512 --seq_len 1024 --max_new_tokens 32 --benchmark
step mixtral_runner 300 python examples/inference/run_mixtral.py --model mixtral-8x7b --layers 2 --max_context_length 256 \
    --seq_len 512 --max_new_tokens 16 --check_accuracy
step dbrx_runner 300 python examples/inference/run_dbrx.py --model dbrx-instruct --layers 2 --max_context_length 256 \
    --seq_len 512 --max_new_tokens 16 --check_accuracy
# 3. training drivers of the other families, one GPU, a few layers
step neox_train 300 python examples/training/tp_dp_family_pretrain.py --family neox --layers 2 --seq_len 2048 --max_steps 5
step bert_train 300 python examples/training/tp_dp_family_pretrain.py --family bert --layers 4 --seq_len 512 --batch_size 8 --max_steps 5
step mixtral_train 300 python examples/training/mixtral_pretrain.py --layers 2 --seq_len 2048 --max_steps 5
step lora_finetune 300 python examples/training/tp_llama_finetune.py --model llama3.2-1b --layers 4 --seq_len 512 \
    --data_dir tools/fixtures/instructions --max_steps 5
# 4. sanitizers over the single-GPU kernel tests (slow: keep the filter narrow when the budget is tight)
SANITIZER_TIMEOUT=900 step sanitizer_memcheck 1000 bash tools/run_sanitizers.sh memcheck "rmsnorm or swiglu or rope or skinny or sampling"
if [ "$NGPU" -ge 2 ]; then
    # 5. multi-GPU: flag-protocol stress, NVLS all-gather variant, TP x PP over NCCL p2p, expert parallelism
    NXD_STRESS=1 step flag_stress_2gpu 400 python -m pytest tests/test_fused_comm_gpu.py -q -m gpu -k flag_protocol_stress -p no:cacheprovider
    NXD_NVLS=1 step nvls_ag_gemm 300 python -m pytest tests/test_fused_comm_gpu.py -q -m gpu -k ag_gemm -p no:cacheprovider
    step pp2_train 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 \
        examples/training/tp_pp_llama_pretrain.py --model llama3.2-1b --layers 4 --tensor_parallel_size 1 \
        --pipeline_parallel_size 2 --num_microbatches 4 --seq_len 2048 --max_steps 5
    step ep2_train 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 \
        examples/training/mixtral_pretrain.py --layers 2 --expert_parallel_size 2 --seq_len 2048 --max_steps 5
fi
echo; cat "$SUMMARY"
