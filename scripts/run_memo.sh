#!/bin/bash
# One-liner kept from the reference's run_memo.sh: start worker 0 of the LR trainer.
PYTHONPATH="$(cd "$(dirname "$0")/.." && pwd)" python -m b200ps.apps.lr2 --job_name=worker --task_index=0 "$@"
