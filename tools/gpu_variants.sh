# Build-and-probe compile-time variants on the GPU box (nvcc is in the image).  Usage: gpu_variants.sh "<flags A>" "<flags B>" ...
mkdir -p gpurun_out
: > gpurun_out/variants.txt
for v in "$@"; do
  echo "== $v" >> gpurun_out/variants.txt
  RTC_NVCC_EXTRA="$v" python -c "from raytracer_challenge_b200 import _build; _build.build_rtc(force=True); _build.build_driver(force=True)" >> gpurun_out/variants.txt 2>&1
  for s in ${RTC_VARIANT_SCENES:-teapot:3840x2160 glass_air_bubble:1920x1080 benchmark:3840x2160 csg_noise:1920x1080}; do
    python tools/perf_probe.py $s 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['scene'], d['ms_kernels'], d['mrays_s'])" >> gpurun_out/variants.txt
  done
done
cat gpurun_out/variants.txt
