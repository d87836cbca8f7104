import sys, os
sys.path.insert(0, os.getcwd())
import litap_b200 as lb
lb._lib.check(lb.lib().litap_init(0))
n=8192
g = lb.DeviceGrid(n+2, n+2); g.synth()
for i in range(4):
    g.slope_gc(5.0); print("gc_ms", lb.timings()["gc_ms"])
g.close()
