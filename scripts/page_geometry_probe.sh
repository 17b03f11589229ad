# C2 (bs64, kv4096, 32/8 heads, bf16, NHD) across page sizes, page orders and loaders: one bench line each
# (value us, roofline frac, isolated us, parity).  Output: gpurun_out/page_geometry.txt
# usage: bash scripts/page_geometry_probe.sh [config ...]   (default: the full table)
out=gpurun_out/page_geometry.txt; : > $out
if [ $# -gt 0 ]; then cfgs=("$@"); else cfgs=("--page-size 16" "--page-size 16 --page-order sequential" "--page-size 16 --loader cpasync" \
           "--page-size 8" "--page-size 8 --page-order sequential" "--page-size 8 --loader cpasync" \
           "--page-size 1" "--page-size 1 --page-order sequential" "--page-size 5" "--page-size 32" "--page-size 64" \
           "--page-size 8 --kv-layout HND" "--page-size 1 --shard-of 8"); fi
for cfg in "${cfgs[@]}"; do
  r=$(python bench.py $cfg --no-extras --no-cpu --steps 100 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],2), round(d['roofline']['frac'],3), round(d['roofline']['kernel_us_isolated'],1), d['parity']['ok'])" 2>&1 | tail -1)
  echo "$cfg: $r" | tee -a $out
done
