"""cProfile of a few bench steps, restricted to this repository's own Python frames (where does the host time go?)."""
import cProfile, pstats, sys, io, os
sys.path.insert(0, os.getcwd())
sys.argv = ["bench.py", "--steps", "5", "--warmup", "3", "--no-cpu-baseline", "--skip-e2e", "--extras", "none", "--no-digest"]
import bench
pr = cProfile.Profile()
pr.enable()
bench.main()
pr.disable()
s = io.StringIO()
st = pstats.Stats(pr, stream=s)
st.sort_stats("tottime").print_stats("repo|numpy|ctypes", 40)
print("\n".join(l for l in s.getvalue().splitlines() if "site-packages/torch" not in l)[:7000])
