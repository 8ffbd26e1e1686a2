"""Linear regression + Adam, then Saver + serving export — the reference's model_export.py.

500 Adam(0.01) steps on 100-point random batches of ``y = x + 20 sin(x/10)`` (1000 points),
prints ``w``, ``b`` and the last loss, then exports to ``<work_dir>/00000003`` with the named
signatures ``inputs: {x}`` / ``outputs: {y}`` (model_export.py:10-11,19-66).
"""
from __future__ import annotations


import numpy as np
import torch

from .. import flags
from ..io.exporter import Exporter, generic_signature
from ..io.saver import Saver
from ..models.linreg import LinearRegression, make_data

FLAGS = flags.FLAGS
flags.DEFINE_string("work_dir", "./model/", "export root (model_export.py:10)")
flags.DEFINE_string("export_version", "0003", "export version, rendered %08d (model_export.py:11)")
flags.DEFINE_string("device", "auto", "cpu | cuda | auto")
flags.DEFINE_integer("export_seed", 0, "RNG seed")

n_samples, learning_rate, batch_size, n_steps = 1000, 0.01, 100, 500


def main(argv=None) -> int:
    dev = FLAGS.device
    if dev == "auto":
        dev = "cuda" if torch.cuda.is_available() else "cpu"
    x_data, y_data = make_data(n_samples)
    xd, yd = torch.from_numpy(x_data).to(dev), torch.from_numpy(y_data).to(dev)
    model = LinearRegression(device=dev, seed=FLAGS.export_seed, scope="test", n_samples=n_samples)
    rng = np.random.RandomState(FLAGS.export_seed)
    loss_val = None
    for _ in range(n_steps):
        indices = torch.from_numpy(rng.choice(n_samples, batch_size)).to(dev)
        loss_val = model.train_step(xd[indices], yd[indices])
    print(model.w.cpu().numpy())
    print(model.b.cpu().numpy())
    print(float(loss_val))

    saver = Saver()
    model_exporter = Exporter(saver)
    model_exporter.init(None, named_graph_signatures={
        "inputs": generic_signature({"x": "Placeholder:0"}),
        "outputs": generic_signature({"y": "test/add:0"}),
    })
    variables = {k: v.detach().cpu().numpy() for k, v in model.variables().items()}
    out = model_exporter.export(FLAGS.work_dir, FLAGS.export_version, variables)
    print("exported to", out)
    return 0


if __name__ == "__main__":
    flags.run(main)
