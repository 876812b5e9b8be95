# BASELINE.json's five configs (and the remaining bins) at their full sizes: device-resident kernel ms, rays, Mrays/s,
# serialised per-class ms, end-to-end ms.  Output: gpurun_out/configs.jsonl
mkdir -p gpurun_out
python tools/perf_probe.py default_world:4096x2160 three_spheres:4096x2160 glass_air_bubble:1920x1080 hexagon:3840x2160 benchmark:3840x2160 \
  teapot:3840x2160 csg_noise:7680x4320:4x4 chapter11_title:4096x2160 chapter12_title:4096x2160 chapter13_title:4096x2160 chapter14_title:4096x2160 cover:4096x4096 \
  > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; tail -3 gpurun_out/configs.err
python -c "
import json
for l in open('gpurun_out/configs.jsonl'):
    d=json.loads(l); print(d['scene'], d['res'], d['spp'], 'ms', d['ms_kernels'], 'Mrays/s', d['mrays_s'], 'rays', d['rays'], 'serial', d['serial_ms'], 'e2e', d['e2e_ms_total'])
"
