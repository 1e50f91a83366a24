mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo "ref rc $?"
SRE_KEEP_REP=none timeout 1500 bash tools/prof_final.sh
