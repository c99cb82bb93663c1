set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1l_tests.txt
cat gpurun_out/r1l_tests.txt
python bench.py > gpurun_out/r1l_bench_D_1gpu.json 2> gpurun_out/r1l_bench_D_1gpu.err || tail -5 gpurun_out/r1l_bench_D_1gpu.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r1l_bench_ref.json 2> gpurun_out/r1l_bench_ref.err || tail -5 gpurun_out/r1l_bench_ref.err
