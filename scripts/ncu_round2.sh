# One `ncu --set full` capture per kernel of libfi_b200.so (north star: "each kernel ships with a
# committed ncu capture"), summarised on the box into profiles/r02_*_summary.md (the .ncu-rep files stay
# in /tmp: ~40 MB each).  Run under gpurun on ONE GPU, after the same commands have exited 0 without ncu.
#   bash scripts/ncu_round2.sh            # everything
#   ONLY=misc bash scripts/ncu_round2.sh  # only the captures whose tag contains $ONLY
set -u
NCU="ncu --set full --clock-control none --import-source on -f"
cap() {  # tag, kernel regex, skip, count, command...
  tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
  case "$tag" in *"${ONLY:-}"*) ;; *) return;; esac
  "$@" > /dev/null 2>&1 || { echo "$tag: command failed WITHOUT ncu, skipped"; return; }
  $NCU -k regex:$rx -s $skip -c $cnt -o /tmp/$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  python scripts/ncu_summary.py /tmp/$tag.ncu-rep $tag "${TRAFFIC[@]:-}" 2>&1 | tail -2
}
B="python bench.py --steps 5 --warmup 3 --no-cpu --no-extras"
TRAFFIC=(--traffic c2); cap r02_decode_c2_bf16 paged_decode 4 2 $B
TRAFFIC=(--traffic c3); cap r02_decode_c3_fp8_tc5 paged_decode 4 2 $B --workload c3
TRAFFIC=()
cap r02_decode_c2_rank_of_8 paged_decode 4 2 $B --shard-of 8
cap r02_decode_c3_rank_of_8 paged_decode 4 2 $B --workload c3 --shard-of 8
cap r02_decode_c3_fp8_page8 paged_decode 4 1 $B --workload c3 --page-size 8
cap r02_decode_c2_bf16_page1_cpasync paged_decode 4 1 $B --page-size 1
cap r02_decode_c2_bf16_page5_cpasync paged_decode 4 1 $B --page-size 5
cap r02_decode_c2_bf16_page8 paged_decode 4 1 $B --page-size 8
cap r02_decode_c2_bf16_rope paged_decode 40 1 python scripts/rope_probe.py ours
cap r02_decode_fp8_d256_tc5 paged_decode_tc5 4 1 python scripts/head_dim_probe.py fp8 256
cap r02_decode_gqa16_rows2 paged_decode 4 1 python scripts/gqa_group_probe.py 128 8
FI_PROBE_ONCE=1 cap r02_kernels_misc 'decode_merge|merge_state|merge_states|append_paged|batch_indices|paged_decode' 0 40 python scripts/misc_kernels_probe.py
case "launches" in *"${ONLY:-}"*)
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file profiles/r02_launches_c2_bf16.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/ncu_launches.log 2>&1;; esac
mkdir -p gpurun_out/profiles && cp profiles/r02_* profiles/decode_kernel_traffic.json gpurun_out/profiles/
ls gpurun_out/profiles
