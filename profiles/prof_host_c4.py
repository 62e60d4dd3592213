"""cProfile of a few configs[3]-shaped steps at 1/8 of the rows (what one of eight ranks sees): where does the host time go?"""
import cProfile, pstats, sys, io, os
sys.path.insert(0, os.getcwd())
sys.argv = ["bench.py", "--workload", "c4", "--rows-per-gpu", "12500000", "--users-per-gpu", "12500", "--steps", "6", "--warmup", "2",
            "--no-cpu-baseline", "--skip-e2e", "--extras", "none", "--no-digest"]
import bench
pr = cProfile.Profile()
pr.enable()
bench.main()
pr.disable()
s = io.StringIO()
st = pstats.Stats(pr, stream=s)
st.sort_stats("cumtime").print_stats("repo", 45)
print("\n".join(l for l in s.getvalue().splitlines() if "site-packages/torch" not in l)[:9000])
