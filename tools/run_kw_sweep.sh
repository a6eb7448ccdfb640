for kw in 256 512 1024; do
  KMLS_KEEP_WORDS=$kw python bench.py --steps 3 --warmup 2 --no-cpu --trace > gpurun_out/bench_kw$kw.json 2>gpurun_out/bench_kw$kw.err
done
