mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q -k "gpu_builder" ) > gpurun_out/gb_pytest.log 2>&1
tail -5 gpurun_out/gb_pytest.log
for w in C2 C3 C4; do for b in lbvh-gpu ploc-gpu; do
  timeout 600 python bench.py --workload $w --builder $b --no-cpu-baseline --steps 10 > gpurun_out/gb_bench_${w}_$b.json 2> gpurun_out/gb_bench_${w}_$b.err
  python -c "
import json;d=json.load(open('gpurun_out/gb_bench_${w}_$b.json'));print('$w $b',d['value'],d['kernel_ms'],d['config']['wide_bvh'])"
done; done
