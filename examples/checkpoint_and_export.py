"""Checkpoint / restore / export with the TF-compatible writers (no TensorFlow involved).

    python examples/checkpoint_and_export.py [--device cpu]

Trains the y = w x + b demo of model_export.py with Adam, saves a Saver-format checkpoint (tensor-bundle V2:
.index + .data-00000-of-00001 + the `checkpoint` state file), restores it, and writes a session_bundle export
directory work_dir/00000003 with the {inputs: x, outputs: y} signature.
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps.io import exporter, saver  # noqa: E402
from b200ps.io.saver import Saver  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--device", default="cuda")
ap.add_argument("--work_dir", default="")
args = ap.parse_args()
work = args.work_dir or tempfile.mkdtemp(prefix="b200ps_export_")

from b200ps.apps import model_export  # noqa: E402
from b200ps import flags  # noqa: E402
flags.FLAGS.work_dir = os.path.join(work, "model")
flags.FLAGS.export_version = "0003"
flags.FLAGS.device = args.device
model_export.main()

sigs, variables = exporter.load_export(os.path.join(work, "model", "00000003"))
print("signatures:", sigs)
print("exported variables:", sorted(variables))
prefix = Saver().save({k: v for k, v in variables.items()}, os.path.join(work, "ckpt", "model.ckpt"), global_step=500)
back = Saver.restore(saver.latest_checkpoint(os.path.join(work, "ckpt")))
print("restored %d variables from %s; w = %.4f, b = %.4f" % (len(back), prefix, float(back["test/weights"].reshape(-1)[0]),
                                                               float(back["test/bias"].reshape(-1)[0])))
