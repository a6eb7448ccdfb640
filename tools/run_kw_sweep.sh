for kw in 64 128 192; do
  KMLS_KEEP_WORDS=$kw python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_kw$kw.json 2>gpurun_out/bench_kw$kw.err
done
