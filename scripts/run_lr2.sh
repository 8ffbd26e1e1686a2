#!/bin/bash
# Launcher for the sparse-LR trainer — the role of the reference's run_lr2.sh, without the
# vendored shflags library or a Hadoop CLI: typed long/short flags are parsed here, train/test
# "directories" are expanded with a local file listing (files > 1000 bytes, comma-joined, as
# run_lr2.sh:39-44 does with `hadoop fs -ls -R | awk`), and one role is started on this host.
#
#   sh scripts/run_lr2.sh -j worker -i 0 -t data/train -T data/test --load_mode all \
#       --learning_rate 0.5 --num_epochs 3 --batch_size 500 --features 4762348
#   sh scripts/run_lr2.sh --run_mode product ...   # cluster submission: all roles via b200ps.launch
set -e
here="$(cd "$(dirname "$0")/.." && pwd)"

job_name=ps; task_index=0
train="data/train"; test="data/test"
run_mode=test; output="output"; load_mode=queue
learning_rate=0.001; num_epochs=120; batch_size=500; features=4762348
num_ps=1; num_workers=2
extra=()

usage() {
  cat <<USAGE
usage: $0 [flags]
  -j, --job_name NAME        job name, ps or worker (default ps)
  -i, --task_index N         task index (default 0)
  -t, --train PATH           train file / directory / comma list
  -T, --test PATH            test file / directory / comma list
      --run_mode MODE        product (all roles on this box) or test (one role) (default test)
      --output PATH          output root
      --load_mode MODE       all or queue (default queue)
      --learning_rate F      (default 0.001)
      --num_epochs N         (default 120)
      --batch_size N         (default 500)
      --features N           (default 4762348)
      --num_ps N / --num_workers N   cluster shape for --run_mode product
  any other --flag[=value] is passed through to b200ps.apps.lr2
USAGE
}

while [ $# -gt 0 ]; do
  case "$1" in
    -h|--help) usage; exit 0 ;;
    -j|--job_name) job_name="$2"; shift 2 ;;        --job_name=*) job_name="${1#*=}"; shift ;;
    -i|--task_index) task_index="$2"; shift 2 ;;    --task_index=*) task_index="${1#*=}"; shift ;;
    -t|--train) train="$2"; shift 2 ;;              --train=*) train="${1#*=}"; shift ;;
    -T|--test) test="$2"; shift 2 ;;                --test=*) test="${1#*=}"; shift ;;
    --run_mode) run_mode="$2"; shift 2 ;;           --run_mode=*) run_mode="${1#*=}"; shift ;;
    --output) output="$2"; shift 2 ;;               --output=*) output="${1#*=}"; shift ;;
    --load_mode) load_mode="$2"; shift 2 ;;         --load_mode=*) load_mode="${1#*=}"; shift ;;
    --learning_rate) learning_rate="$2"; shift 2 ;; --learning_rate=*) learning_rate="${1#*=}"; shift ;;
    --num_epochs) num_epochs="$2"; shift 2 ;;       --num_epochs=*) num_epochs="${1#*=}"; shift ;;
    --batch_size) batch_size="$2"; shift 2 ;;       --batch_size=*) batch_size="${1#*=}"; shift ;;
    --features) features="$2"; shift 2 ;;           --features=*) features="${1#*=}"; shift ;;
    --num_ps) num_ps="$2"; shift 2 ;;               --num_ps=*) num_ps="${1#*=}"; shift ;;
    --num_workers) num_workers="$2"; shift 2 ;;     --num_workers=*) num_workers="${1#*=}"; shift ;;
    *) extra+=("$1"); shift ;;
  esac
done
case "$task_index$num_epochs$batch_size$features" in *[!0-9]*) echo "integer flag expected" >&2; exit 1 ;; esac

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
