#!/bin/bash
# GPT-NeoX 20B: TP8 x DP; reference tp_dp_gpt_neox_20b_hf_pretrain.sh
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
GBS=${GBS:-256}
MBS=${MBS:-1}
DATA_PATH=${DATA_PATH:-}
torchrun $DISTRIBUTED_ARGS $EXAMPLES/tp_dp_family_pretrain.py --family neox \
    --model_path $CONFIGS/gpt_neox_20b ${DATA_PATH:+--data_dir $DATA_PATH} \
    --tensor_parallel_size ${TP_DEGREE:-8} --seq_len ${SEQ_LEN:-2048} --lr ${LR:-9.7e-5} --warmup_steps 2000 \
    --max_steps ${STEPS_OVERRIDE:-${TOTAL_STEPS:-10000}} \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_gpt_neox_20b} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --loading_step latest_if_exists --tensorboard 1 --output_dir ${OUTPUT_DIR:-./output_gpt_neox_20b} \
    "$@" 2>&1 | tee "$LOG_DIR/log_gpt_neox_20b-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
