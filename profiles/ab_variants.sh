set -e
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-migration 2>gpurun_out/err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['kernels']['service_pass_ms'])"; }
echo "== base"; run
for V in "-DUVMTIER_X_LATECOLLECT" "-DUVMTIER_X_ASMLOADS" "-DUVMTIER_X_LATECOLLECT -DUVMTIER_X_ASMLOADS"; do
  (cd open-gpu-kernel-modules_b200/csrc && touch service.cu && make -j8 EXTRA="$V" 2>&1 | grep -E "error" || true)
  echo "== $V"; run
done
