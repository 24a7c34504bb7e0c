import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
bench.EQUIL_STEPS = 3000
state, L = bench.make_state("lj64k")
s = bench.new_system(state, L)
s.run(2000); s.run(2000)
host = s.particles()
for it in range(4):
    t0 = time.perf_counter(); s.set_particles(host["pos"], host["vel"], host["image"]); t1 = time.perf_counter()
    el, ms = s.run(2000); t2 = time.perf_counter()
    host = s.particles(); t3 = time.perf_counter()
    th = s.thermo(); t4 = time.perf_counter()
    print("set %.1f ms | run wall %.1f ms (loop %.1f ms, gpu %.1f ms) | get %.1f ms | thermo %.1f ms | stats %s" % (
        1e3*(t1-t0), 1e3*(t2-t1), 1e3*el, ms, 1e3*(t3-t2), 1e3*(t4-t3), s.stats()), flush=True)
