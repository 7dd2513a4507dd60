"""Host-side profile of the training bench on the tiny workload: who calls what (pstats callers of the hot built-ins)."""
import cProfile, pstats, sys, io, runpy
sys.argv_callers = sys.argv[1:] or ["_cuda_getDeviceCount"]
sys.argv = ["bench.py", "--mode", "train", "--workload", "tiny", "--steps", "200", "--warmup", "5", "--no-cpu-baseline"]
pr = cProfile.Profile()
pr.enable()
try:
    runpy.run_path("bench.py", run_name="__main__")
except SystemExit:
    pass
pr.disable()
s = io.StringIO()
ps = pstats.Stats(pr, stream=s)
for name in sys.argv_callers:
    ps.print_callers(name)
print(s.getvalue()[:20000], file=sys.stderr)
