"""cProfile of bench.py's cfg3 loop (where the host time of rate_best goes)."""
import cProfile, pstats, sys, io
sys.path.insert(0, '.')
import bench
sys.argv = ['bench.py', '--workload', 'cfg3', '--steps', '2', '--warmup', '1', '--no-cpu']
pr = cProfile.Profile()
pr.enable()
bench.main()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(28)
print(s.getvalue()[:6000], file=sys.stderr)
