#!/bin/bash
# A/B of library builds (tools/build_variant.sh): one short bench run per variant, same box.
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense > gpurun_out/var_base.json 2> gpurun_out/var_base.err
for v in "$@"; do
  KMLS_LIB=$PWD/kubernetes-machine-learning-server_b200/variants/libkmls_$v.so python bench.py --steps 3 --warmup 2 --no-cpu --no-dense > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
done
