#!/bin/bash
# BERT-large phase-2 style pre-training: TP2 x DP; reference tp_dp_bert_large_hf_pretrain_hdf5.py
set -e
source "$(dirname "${BASH_SOURCE[0]}")/_launch_common.sh"
GBS=${GBS:-256}
MBS=${MBS:-1}
DATA_PATH=${DATA_PATH:-}
torchrun $DISTRIBUTED_ARGS $EXAMPLES/tp_dp_family_pretrain.py --family bert \
    --model_path $CONFIGS/bert_large ${DATA_PATH:+--data_dir $DATA_PATH} \
    --tensor_parallel_size ${TP_DEGREE:-2} --seq_len ${SEQ_LEN:-512} --batch_size ${MBS:-16} --lr ${LR:-4e-4} --warmup_steps 2000 \
    --max_steps ${STEPS_OVERRIDE:-${TOTAL_STEPS:-10000}} \
    --checkpoint_dir ${CHECKPOINT_DIR:-./ckpt_bert_large} --checkpoint_freq ${CKPT_FREQ_OVERRIDE:-${CHECKPOINT_FREQ:-1000}} \
    --loading_step latest_if_exists --tensorboard 1 --output_dir ${OUTPUT_DIR:-./output_bert_large} \
    "$@" 2>&1 | tee "$LOG_DIR/log_bert_large-$NODE_RANK.log"
exit ${PIPESTATUS[0]}
