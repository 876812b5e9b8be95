"""Image::par_render of the host API from ONE process over 1..N GPUs: wall time per call, slowest device's kernel time.
Usage: one_process_probe.py [scene:WxH[:SXxSY]] [n_devices list, e.g. 1,2] [reps]"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
case = sys.argv[1] if len(sys.argv) > 1 else "teapot:3840x2160"
counts = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1,2").split(",")]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 6
name, res, *rest = case.split(":")
w, h = map(int, res.split("x"))
spp = tuple(map(int, rest[0].split("x"))) if rest else (1, 1)
drv = R.load_driver()
sess = R.Session(name, w, h, device=0)
err = C.create_string_buffer(1024)
probe, phases, st = (C.c_double * 3)(), (C.c_double * 2)(), R.Stats()
opts = R.Opts(spp=spp)
for graphs in (1, 0):
    R.set_tuning(no_graphs=0 if graphs else 1)
    for n in counts:
        times = []
        for _ in range(reps):
            t0 = time.perf_counter()
            rc = drv.drv_par_render_step(sess.handle, 0, n, C.byref(opts), probe, C.byref(st), phases, err, 1024)
            if rc:
                raise SystemExit(err.value.decode())
            times.append((time.perf_counter() - t0) * 1e3)
        print(f"graphs={graphs} devices={n}: wall best {min(times):.2f} ms, median {sorted(times)[len(times)//2]:.2f} ms, slowest device kernels {st.ms_kernels:.2f} ms, flatten {phases[0]:.2f} ms, render call {st.ms_total:.2f} ms", flush=True)
