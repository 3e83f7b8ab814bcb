# development: runs tools/probe.py on the central slab of the bench grid with the diagnostic libraries given as arguments
mkdir -p gpurun_out
for b in "$@"; do
  if [ "$b" = 0 ]; then lib=critic2_b200/libc2gpu.so; else lib=critic2_b200/libc2gpu_diag$b.so; fi
  echo "=== C2_WS_DIAG=$b"
  C2GPU_LIB=$PWD/$lib timeout 300 python tools/probe.py --cfgs 6 --z0 196 --z1 204 2>&1 | grep -v "^$" | tail -7
done
