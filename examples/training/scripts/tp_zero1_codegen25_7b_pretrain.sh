#!/bin/bash
# CodeGen2.5 7B: TP8 + ZeRO-1; reference examples/training/codegen25/tp_zero1_codegen25_7b_hf_pretrain.sh
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
GBS=${GBS:-256}
MBS=${MBS:-1}
DATA_PATH=${DATA_PATH:-}
torchrun $DISTRIBUTED_ARGS $EXAMPLES/tp_dp_family_pretrain.py --family codegen \
    --model_path $CONFIGS/codegen25_7b ${DATA_PATH:+--data_dir $DATA_PATH} \
    --tensor_parallel_size ${TP_DEGREE:-8} --seq_len ${SEQ_LEN:-2048} --lr ${LR:-3e-5} --warmup_steps 100 \
    --max_steps ${STEPS_OVERRIDE:-${TOTAL_STEPS:-10000}} \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_codegen25_7b} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --loading_step latest_if_exists --tensorboard 1 --output_dir ${OUTPUT_DIR:-./output_codegen25_7b} \
    "$@" 2>&1 | tee "$LOG_DIR/log_codegen25_7b-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
