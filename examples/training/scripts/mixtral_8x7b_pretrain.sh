#!/bin/bash
# Mixtral 8x7B: TP x EP (experts sharded over EP, EP-aware ZeRO-1); reference examples/training/mixtral/mixtral_pretrain/run_mixtral_pretrain_ptl.sh
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
GBS=${GBS:-256}
MBS=${MBS:-1}
DATA_PATH=${DATA_PATH:-}
torchrun $DISTRIBUTED_ARGS $EXAMPLES/mixtral_pretrain.py \
    --model_path $CONFIGS/mixtral_8x7b ${DATA_PATH:+--data_dir $DATA_PATH} \
    --tensor_parallel_size ${TP_DEGREE:-4} --expert_parallel_size ${EP_DEGREE:-2} --seq_len ${SEQ_LEN:-4096} --lr ${LR:-1e-4} --warmup_steps 100 --capacity_factor 2.0 \
    --max_steps ${STEPS_OVERRIDE:-${TOTAL_STEPS:-10000}} \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_mixtral_8x7b} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --loading_step latest_if_exists --tensorboard 1 --output_dir ${OUTPUT_DIR:-./output_mixtral_8x7b} \
    "$@" 2>&1 | tee "$LOG_DIR/log_mixtral_8x7b-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
