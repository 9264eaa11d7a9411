import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import test_gpu_parity as t
which = sys.argv[1]
if which == 'a': t.run_gbt(30000, [64] * 40, seed=18, n_trees=1, max_depth=7, weighted=True)
if which == 'b': t.run_gbt(30000, [64] * 40, seed=18, n_trees=1, max_depth=7, weighted=True, flags=2)
if which == 'c': t.run_gbt(30000, [64] * 40, seed=19, n_trees=2, max_depth=6, bagging=True, feature_subsets=True)
print('ok', which)
