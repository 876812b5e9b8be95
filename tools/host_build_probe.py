import time, ctypes as C, os, sys
import raytracer_challenge_b200 as R
lib = R.load_driver(); rtc = R.load_rtc()
h = C.c_void_p(); st = (C.c_int64*16)(); err = C.create_string_buffer(1024)
rc = lib.drv_open(b"teapot", R.DATA_DIR.encode(), 3840, 2160, -1, C.byref(h), st, err, 1024)
class Desc(C.Structure): _fields_=[("x", C.c_char*512)]
lib.drv_desc_tmp.restype = C.c_void_p
lib.drv_desc_tmp.argtypes=[C.c_void_p, C.c_void_p]
d = Desc()
pd = lib.drv_desc_tmp(h, C.byref(d))
best=1e9
for i in range(20):
    t=time.perf_counter(); rtc.rtc_scene_validate(C.c_void_p(pd), st); best=min(best,(time.perf_counter()-t)*1e3)
print("best validate ms", round(best,3), "nodes", st[1], "depth", st[2])
