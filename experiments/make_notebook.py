"""Build ``experiments/experiments.ipynb`` — the counterpart of the reference's notebook.

The cells are executed here (plain ``exec`` with captured stdout, no Jupyter needed) and stored
with their outputs, so the notebook can be read on its own and re-run with any Jupyter front end.
Figures are the PNG files written by ``experiments.py`` (its built-in writer; matplotlib is used
when it is installed).

    python experiments/make_notebook.py
"""
import base64
import contextlib
import io
import json
import pathlib
import sys

HERE = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parent))

CELLS = [
    ("markdown", """# mpitree_b200 experiments

Counterpart of the reference's `experiments.ipynb` (parallel tree on 4 ranks, decision boundaries,
serial tree, timing table).  Everything here runs on the host engine (`device="cpu"`), so the
notebook works on any machine; on a B200 box drop the `device` argument (or pass `"cuda"`) and
launch multi-GPU runs with `torchrun`, one rank per GPU.""", None),
    ("markdown", """## Parallel tree, 4 ranks, iris[:, :2], `max_depth=3`

The reference runs this cell with `mpirun -n 4`.  Here the four ranks are threads of this process
(`run_threaded`); `torchrun --nproc-per-node 4` with the same estimator code uses real processes.""", None),
    ("code", """from sklearn.datasets import load_iris
from mpitree_b200.parallel.comm import run_threaded
from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

iris = load_iris()
X, y = iris.data[:, :2], iris.target

def rank_main(comm):
    clf = ParallelDecisionTreeClassifier(max_depth=3, comm=comm, device="cpu").fit(X, y)
    if not comm.rank:
        print(clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names))

_ = run_threaded(4, rank_main)""", None),
    ("markdown", "## Decision boundaries, `max_depth` 2 and 5", None),
    ("code", """import experiments  # experiments/experiments.py
experiments.cell3()""", ["decision_boundary_depth2.png", "decision_boundary_depth5.png"]),
    ("markdown", "## Serial tree, `max_depth=5`, `precision=1`", None),
    ("code", """clf = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
print(clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names, precision=1))""", None),
    ("markdown", """## Fit time vs. number of samples

`n` distinct classes, one feature, unlimited depth, as in the reference's timing cell.  The
parallel rows are measured with real processes by `parallel_timing.py` (one `torchrun` launch
per `k`); all ranks receive the full data, like the reference.""", None),
    ("code", """x_dim, rows = experiments.cell5()
experiments.plots(x_dim, rows)""", ["timing.png"]),
    ("markdown", """A larger replicated problem (3000 x 8 continuous features, unlimited depth), same processes: every
rank builds the top of the tree redundantly, the subtrees below `n / k` rows are dealt out, one
all-gather returns them (`parallel/scheduler.py::fit_flat_tree_replicated`).""", None),
    ("code", """import json, pathlib
for k in (2, 5, 8):
    f = pathlib.Path(experiments.HERE) / f"timing_large_k{k}.json"
    if f.exists():
        r = json.loads(f.read_text())
        print(f"k={k}: serial {r['serial_ms']:.0f} ms, parallel {r['parallel_ms']:.0f} ms "
              f"({r['serial_ms'] / r['parallel_ms']:.1f}x)")""", None),
]


def main():
    ns = {"__name__": "__notebook__"}
    cells = []
    count = 0
    for kind, src, images in CELLS:
        if kind == "markdown":
            cells.append({"cell_type": "markdown", "metadata": {}, "source": src.splitlines(keepends=True)})
            continue
        count += 1
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            exec(compile(src, f"<cell {count}>", "exec"), ns)  # noqa: S102 - our own cells
        outputs = []
        if buf.getvalue():
            outputs.append({"output_type": "stream", "name": "stdout", "text": buf.getvalue().splitlines(keepends=True)})
        for name in images or []:
            path = HERE / name
            if path.exists():
                outputs.append({"output_type": "display_data", "metadata": {},
                                "data": {"image/png": base64.b64encode(path.read_bytes()).decode(),
                                         "text/plain": [f"<{name}>"]}})
        cells.append({"cell_type": "code", "execution_count": count, "metadata": {}, "outputs": outputs,
                      "source": src.splitlines(keepends=True)})
    nb = {
        "cells": cells,
        "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                     "language_info": {"name": "python"}},
        "nbformat": 4,
        "nbformat_minor": 5,
    }
    out = HERE / "experiments.ipynb"
    out.write_text(json.dumps(nb, indent=1) + "\n")
    print(f"wrote {out} ({len(cells)} cells)")


if __name__ == "__main__":
    main()
