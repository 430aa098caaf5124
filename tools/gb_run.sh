mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q -k "gpu_builder" ) > gpurun_out/gb_pytest.log 2>&1
tail -15 gpurun_out/gb_pytest.log
for b in sah-host lbvh-gpu; do
  timeout 300 python bench.py --builder $b --no-cpu-baseline --steps 20 > gpurun_out/gb_bench_C2_$b.json 2> gpurun_out/gb_bench_C2_$b.err
  python -c "
import json;d=json.load(open('gpurun_out/gb_bench_C2_$b.json'));print('$b',d['value'],d['kernel_ms'],d['config']['wide_bvh'])"
done
