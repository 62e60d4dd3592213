import cProfile, pstats, sys, io, os
sys.path.insert(0, os.getcwd())
sys.argv = ["bench.py", "--steps", "3", "--warmup", "2", "--no-cpu-baseline", "--skip-e2e"]
import bench
pr = cProfile.Profile()
pr.enable()
bench.main()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print(s.getvalue()[:9000])
