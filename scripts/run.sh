#!/bin/bash
# File-list builder — the role of the reference's run.sh: list the train/test inputs and write
# one path per line to data/{train,test}_file_list (run.sh:25-41 does it with `hadoop fs -ls`
# and uploads the list back to HDFS; here the lists are local files that
# `--train_file_list/--test_file_list` of b200ps.apps.lr2 read).
set -e
data="data"; mkdir -p "$data"
train="data/train"; test="data/test"
train_file_list="$data/train_file_list"; test_file_list="$data/test_file_list"
while [ $# -gt 0 ]; do
  case "$1" in
    -t|--train) train="$2"; shift 2 ;;  --train=*) train="${1#*=}"; shift ;;
    -T|--test) test="$2"; shift 2 ;;    --test=*) test="${1#*=}"; shift ;;
    --train_file_list) train_file_list="$2"; shift 2 ;;  --train_file_list=*) train_file_list="${1#*=}"; shift ;;
    --test_file_list) test_file_list="$2"; shift 2 ;;    --test_file_list=*) test_file_list="${1#*=}"; shift ;;
    -h|--help) echo "usage: $0 [-t train_dir] [-T test_dir] [--train_file_list F] [--test_file_list F]"; exit 0 ;;
    *) echo "unknown flag $1" >&2; exit 1 ;;
  esac
done
set -x
get_file_list() { rm -f "$2"; find "$1" -type f | sort > "$2"; }
get_file_list "$train" "$train_file_list"
get_file_list "$test" "$test_file_list"
