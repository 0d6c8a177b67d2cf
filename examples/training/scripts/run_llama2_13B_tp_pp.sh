#!/bin/bash
# llama2_13b: tensor parallel (8) x pipeline parallel (4, 1F1B) + ZeRO-1 over data parallel, selective activation checkpointing,
# meta-device initialisation.  Counterpart of the reference's tp_pp_llama_hf_pretrain launch scripts.
# 70B needs several nodes: submit with sbatch/srun so that SLURM_NTASKS / SLURM_NODEID are set.
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
TP_DEGREE=${TP_DEGREE:-8}
PP_DEGREE=${PP_DEGREE:-4}
GBS=${GBS:-1024}
SEQ_LEN=${SEQ_LEN:-4096}
DP=$((WORLD_GPUS / TP_DEGREE / PP_DEGREE))
[ "$DP" -ge 1 ] || { echo "$WORLD_GPUS GPUs cannot hold TP=$TP_DEGREE x PP=$PP_DEGREE"; exit 1; }
NUM_MICROBATCHES=$((GBS / DP))
MAX_STEPS=${STEPS_OVERRIDE:-${MAX_STEPS:-30000}}
DATA_PATH=${DATA_PATH:-}
echo "TP=$TP_DEGREE PP=$PP_DEGREE DP=$DP GBS=$GBS microbatches=$NUM_MICROBATCHES SEQ_LEN=$SEQ_LEN"
torchrun $DISTRIBUTED_ARGS $EXAMPLES/tp_pp_llama_pretrain.py \
    --training_config $CONFIGS/llama2_13b \
    ${DATA_PATH:+--training_dir $DATA_PATH} \
    --tensor_parallel_size $TP_DEGREE --pipeline_parallel_size $PP_DEGREE \
    --num_microbatches $NUM_MICROBATCHES --micro_batch_size 1 --seq_len $SEQ_LEN \
    --max_steps $MAX_STEPS --lr ${LR:-3e-4} --min_lr 1e-6 --beta1 0.9 --beta2 0.95 --weight_decay 0.1 \
    --warmup_steps ${WARMUP_STEPS:-2000} --constant_steps 0 \
    --use_zero1_optimizer 1 --use_selective_checkpoint 1 --activation_checkpointing 0 --use_meta_device_init 1 \
    --kv_replicator ${KV_REPLICATOR:-1} --save_load_xser 1 \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_llama2_13b} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --num_kept_checkpoint -1 --loading_step latest_if_exists --tb_dir ${TB_DIR:-./tb_llama2_13b} \
    "$@" 2>&1 | tee "$LOG_DIR/log_llama2_13b-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
