"""Drop-in ``mpitree`` import path and the bench.py JSON contract (reference arm; CPU only)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mpitree_import_alias_runs_notebook_cell():
    code = (
        "from sklearn.datasets import load_iris\n"
        "from mpitree.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier\n"
        "from mpitree.tree._base import Node, BranchType\n"
        "iris = load_iris(); X, y = iris.data[:, :2], iris.target\n"
        "clf = ParallelDecisionTreeClassifier(max_depth=3, device='cpu').fit(X, y)\n"
        "assert not clf.WORLD_RANK\n"
        "print(clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names).splitlines()[0])\n"
    )
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "compat"), ROOT]))
    res = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    assert res.stdout.strip() == "┌── sepal length (cm)"


def test_bench_reference_arm_prints_one_json_line():
    env = dict(os.environ, MTB_REF_BUDGET_S="1")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--features", "8", "--depth", "3"],
                         env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference"
    if "unavailable" not in rec:
        for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config"):
            assert key in rec
        assert rec["value"] > 0 and rec["config"]["rows"] >= 60


def test_notebook_holds_the_reference_goldens():
    """experiments/experiments.ipynb is stored with outputs; its two trees are the reference's goldens."""
    from .goldens import IRIS2_DEPTH3, IRIS2_DEPTH5_P1

    nb = json.load(open(os.path.join(ROOT, "experiments", "experiments.ipynb"), encoding="utf-8"))
    assert nb["nbformat"] == 4
    texts = ["".join(o["text"]) for c in nb["cells"] if c["cell_type"] == "code"
             for o in c["outputs"] if o["output_type"] == "stream"]
    assert IRIS2_DEPTH3.strip() in [t.strip() for t in texts]
    assert IRIS2_DEPTH5_P1.strip() in [t.strip() for t in texts]
    images = [o for c in nb["cells"] if c["cell_type"] == "code" for o in c["outputs"] if o["output_type"] == "display_data"]
    assert len(images) == 3  # two decision boundaries and the timing plot
