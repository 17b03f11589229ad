# One GPU-box pass: tests, bench lines, launch list and ncu --set full captures, summarised on the
# box into gpurun_out/profiles/ (the .ncu-rep files are ~40 MB each; only the C2 one travels back).
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; tail -2 gpurun_out/t.log
python bench.py > gpurun_out/b_c2.log 2>&1; tail -1 gpurun_out/b_c2.log
python bench.py --workload c3 --no-cpu > gpurun_out/b_c3.log 2>&1; tail -1 gpurun_out/b_c3.log
python scripts/rope_probe.py ours | tee gpurun_out/rope.log
TAG=${TAG:-r01c}
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches_c2_bf16.csv python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 4 -c 2 -f -o gpurun_out/${TAG}_c2 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 4 -c 2 -f -o /tmp/${TAG}_c3 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu > gpurun_out/ncu_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 40 -c 2 -f -o /tmp/${TAG}_rope python scripts/rope_probe.py ours > gpurun_out/ncu_rope.log 2>&1
python scripts/ncu_summary.py gpurun_out/${TAG}_c2.ncu-rep ${TAG}_decode_c2_bf16 --traffic c2
python scripts/ncu_summary.py /tmp/${TAG}_c3.ncu-rep ${TAG}_decode_c3_fp8_tc5 --traffic c3
python scripts/ncu_summary.py /tmp/${TAG}_rope.ncu-rep ${TAG}_decode_c2_bf16_rope
mkdir -p gpurun_out/profiles && cp profiles/${TAG}_* profiles/decode_kernel_traffic.json gpurun_out/profiles/
ls -la gpurun_out gpurun_out/profiles
