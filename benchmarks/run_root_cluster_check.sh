export PYTHONPATH=.
mkdir -p gpurun_out
timeout 120 python - <<'PY'
import torch, numpy as np, time
from mpitree_b200.tree import DecisionTreeClassifier
from bench import synth_rows
import os
dev=torch.device("cuda",0)
for F in (64, 32, 128):
  for n in (1000, 70001, 1_000_000):
    X,y=synth_rows(torch,dev,0,n,F,10)
    os.environ["MTB_ROOT_CLUSTER"]="1"
    a=DecisionTreeClassifier(max_depth=6,device="cuda").fit(X,y)
    os.environ["MTB_ROOT_CLUSTER"]="0"
    b=DecisionTreeClassifier(max_depth=6,device="cuda").fit(X,y)
    fa,fb=a.flat_tree_,b.flat_tree_
    print(F, n, fa.n_nodes, fb.n_nodes, np.array_equal(fa.feature,fb.feature), np.array_equal(fa.counts,fb.counts), np.array_equal(fa.bin,fb.bin))
print("root cluster ok")
PY
echo "exit $?"
for mc in 1 0; do
echo "--- MTB_ROOT_MULTICAST=$mc"
MTB_ROOT_MULTICAST=$mc timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 --root-cluster --depth 2 2>&1 | grep -v -i warn | tail -3 | cut -c1-250
done
