#!/bin/bash
# File-list builder — the role of the reference's run.sh: list the train/test inputs and write
# one path per line to data/{train,test}_file_list (run.sh:25-41 does it with `hadoop fs -ls`
# and uploads the list back to HDFS; here the lists are local files that
# `--train_file_list/--test_file_list` of b200ps.apps.lr2 read).
set -e
here="$(cd "$(dirname "$0")/.." && pwd)"
. "$here/scripts/bashflags"
DEFINE_string train "data/train" "directory of training files" t
DEFINE_string test "data/test" "directory of test files" T
DEFINE_string train_file_list "data/train_file_list" "where the train list is written"
DEFINE_string test_file_list "data/test_file_list" "where the test list is written"
FLAGS "$@" || { rc=$?; [ $rc -eq ${FLAGS_FALSE} ] && exit 0; exit $rc; }
train=${FLAGS_train}; test=${FLAGS_test}
train_file_list=${FLAGS_train_file_list}; test_file_list=${FLAGS_test_file_list}
mkdir -p "$(dirname "$train_file_list")" "$(dirname "$test_file_list")"
set -x
get_file_list() { rm -f "$2"; find "$1" -type f | sort > "$2"; }
get_file_list "$train" "$train_file_list"
get_file_list "$test" "$test_file_list"
