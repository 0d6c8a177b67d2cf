#!/bin/bash
# llama3.2_1b: tensor parallel (8) + sequence parallel + ZeRO-1 over the remaining data-parallel ranks.
# Counterpart of the reference's tp_zero1_llama_hf_pretrain launch scripts (same global batch / sequence / schedule).
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
TP_DEGREE=${TP_DEGREE:-8}
GBS=${GBS:-256}
MBS=${MBS:-1}
SEQ_LEN=${SEQ_LEN:-8192}
LR=${LR:-3e-4}
WARMUP_STEPS=${WARMUP_STEPS:-100}
TOTAL_STEPS=${STEPS_OVERRIDE:-${TOTAL_STEPS:-10000}}
DATA_PATH=${DATA_PATH:-}                       # tokenised dataset (get_dataset.py); synthetic tokens when empty
DP=$((WORLD_GPUS / TP_DEGREE))
ACC_STEPS=$((GBS / MBS / DP))
[ "$ACC_STEPS" -ge 1 ] || { echo "GBS=$GBS is smaller than MBS*DP=$((MBS * DP))"; exit 1; }
echo "TP=$TP_DEGREE DP=$DP GBS=$GBS MBS=$MBS ACC_STEPS=$ACC_STEPS SEQ_LEN=$SEQ_LEN LR=$LR STEPS=$TOTAL_STEPS"
torchrun $DISTRIBUTED_ARGS $EXAMPLES/tp_zero1_llama_pretrain.py \
    --model_path $CONFIGS/llama3.2_1b \
    ${DATA_PATH:+--data_dir $DATA_PATH} \
    --tensor_parallel_size $TP_DEGREE --use_sequence_parallel 1 --selective_checkpoint 1 \
    --batch_size $MBS --grad_accum_usteps $ACC_STEPS --seq_len $SEQ_LEN \
    --max_steps $TOTAL_STEPS --warmup_steps $WARMUP_STEPS --lr $LR \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_llama3.2_1b} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --loading_step latest_if_exists --tensorboard 1 --output_dir ${OUTPUT_DIR:-./output_llama3.2_1b} \
    "$@" 2>&1 | tee "$LOG_DIR/log_llama3.2_1b-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
