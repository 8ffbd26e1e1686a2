#!/bin/bash
# Launcher for the sparse-LR trainer — the role of the reference's run_lr2.sh, without a Hadoop CLI:
# typed long/short flags come from scripts/bashflags (this tree's shflags replacement), train/test
# "directories" are expanded with a local file listing (files > 1000 bytes, comma-joined, as
# run_lr2.sh:39-44 does with `hadoop fs -ls -R | awk`), and one role is started on this host.
#
#   sh scripts/run_lr2.sh -j worker -i 0 -t data/train -T data/test --load_mode all \
#       --learning_rate 0.5 --num_epochs 3 --batch_size 500 --features 4762348
#   sh scripts/run_lr2.sh --run_mode product ...   # cluster submission: all roles via b200ps.launch
set -e
here="$(cd "$(dirname "$0")/.." && pwd)"

. "$here/scripts/bashflags"
FLAGS_HELP="usage: $0 [flags] [-- extra b200ps.apps.lr2 flags]
  --run_mode product starts every role on this box through b200ps.launch; test starts one role.
  Unknown --flags are passed through to b200ps.apps.lr2."
FLAGS_PASSTHROUGH=1
DEFINE_string  job_name ps "job name: ps or worker" j
DEFINE_integer task_index 0 "task index" i
DEFINE_string  train "data/train" "train file / directory / comma list" t
DEFINE_string  test "data/test" "test file / directory / comma list" T
DEFINE_string  run_mode test "product (all roles via b200ps.launch) or test (one role)"
DEFINE_string  output output "output root"
DEFINE_string  load_mode queue "all or queue"
DEFINE_float   learning_rate 0.001 "SGD learning rate"
DEFINE_integer num_epochs 120 "epochs"
DEFINE_integer batch_size 500 "examples per step"
DEFINE_integer features 4762348 "feature-id space"
DEFINE_integer num_ps 1 "ps tasks for --run_mode product"
DEFINE_integer num_workers 2 "worker tasks for --run_mode product"
FLAGS "$@" || { rc=$?; [ $rc -eq ${FLAGS_FALSE} ] && exit 0; exit $rc; }
eval "extra=(${FLAGS_ARGV})"
job_name=${FLAGS_job_name}; task_index=${FLAGS_task_index}; train=${FLAGS_train}; test=${FLAGS_test}
run_mode=${FLAGS_run_mode}; output=${FLAGS_output}; load_mode=${FLAGS_load_mode}
learning_rate=${FLAGS_learning_rate}; num_epochs=${FLAGS_num_epochs}; batch_size=${FLAGS_batch_size}
features=${FLAGS_features}; num_ps=${FLAGS_num_ps}; num_workers=${FLAGS_num_workers}

get_file_list() {   # comma-joined files larger than 1000 bytes under a path (or the path itself)
  local p="$1"
  if [ "$p" = "synthetic" ] || [ -f "$p" ] || echo "$p" | grep -q ","; then echo "$p"; return; fi
  find "$p" -type f -size +1000c 2>/dev/null | sort | tr '\n' ',' | sed 's/,$//'
}
train_data=$(get_file_list "$train")
test_data=$(get_file_list "$test")
echo "train data: ${train_data}"
echo "test data: ${test_data}"
echo "run mode: ${run_mode}"
echo "load mode: ${load_mode}"
set -x
common=(--train="${train_data}" --test="${test_data}" --mode="${load_mode}" --learning_rate="${learning_rate}"
        --num_epochs="${num_epochs}" --batch_size="${batch_size}" --features="${features}")
if [ "${run_mode}" = "product" ]; then
  PYTHONPATH="$here" exec python -m b200ps.launch --num_ps "${num_ps}" --num_workers "${num_workers}" \
      --workdir "${output}" lr2 -- "${common[@]}" "${extra[@]}"
else
  PYTHONPATH="$here" exec python -m b200ps.apps.lr2 --job_name="${job_name}" --task_index="${task_index}" \
      "${common[@]}" "${extra[@]}"
fi
