#!/usr/bin/env bash
# Launcher for the sm_100a path with the command line of the reference's launcher (src/demo.sh:1-37):
#
#   demo.sh train [extra flags]          NAML, category + subcategory views, frozen title table, 5 epochs of batch 32
#   demo.sh test <checkpoint> [flags]    evaluation with user_log_mask=True, batch 128 (train/test differ here)
#   demo.sh embeddings                   check that the title tables of both data directories load
#
# Every setting can be overridden from the environment (model, model_dir, nGPU, epochs, batch_size,
# train_data_dir, test_data_dir, prepare); anything after the sub-command's own arguments is handed to
# `python -m newsrecommendation_b200.main` unchanged.
set -euo pipefail

usage() { echo "Please select train or test mode." >&2; }

sub=${1:-}
[ -n "${sub}" ] || { usage; exit 2; }
shift

repo_root=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
run_main() { PYTHONPATH="${repo_root}${PYTHONPATH:+:${PYTHONPATH}}" python -u -m newsrecommendation_b200.main "$@"; }

: "${model:=NAML}"
: "${model_dir:=../model/${model}}"
: "${nGPU:=1}"
: "${prepare:=False}"
: "${train_data_dir:=../data/MINDsmall_train}"
: "${test_data_dir:=../data/MINDsmall_dev}"

# what both sub-commands share: the model family, its views, the frozen table, the device count
common=(--model "${model}" --model_dir "${model_dir}" --nGPU "${nGPU}" --enable_gpu True --prepare "${prepare}"
        --use_category True --use_subcategory True --freeze_embedding True)

case "${sub}" in
  train)
    run_main --mode train "${common[@]}" --train_data_dir "${train_data_dir}" \
             --epochs "${epochs:-5}" --batch_size "${batch_size:-32}" --lr 0.0003 --user_log_mask False "$@"
    ;;
  test)
    ckpt=${1:-}
    [ -n "${ckpt}" ] || { echo "usage: demo.sh test <checkpoint name in ${model_dir}>" >&2; exit 2; }
    shift
    run_main --mode test "${common[@]}" --test_data_dir "${test_data_dir}" \
             --batch_size "${batch_size:-128}" --user_log_mask True --load_ckpt_name "${ckpt}" "$@"
    ;;
  embeddings)
    run_main --mode read_embeddings --train_data_dir "${train_data_dir}" --test_data_dir "${test_data_dir}" "$@"
    ;;
  *)
    usage
    exit 2
    ;;
esac
