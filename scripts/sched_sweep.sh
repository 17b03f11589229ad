for cfg in "100 50 4 4" "85 50 2 4" "85 50 4 6" "70 50 4 6" "90 100 2 4"; do set -- $cfg
 for a in "c2 8" "c3 8"; do set -- $cfg; w=${a% *}; s=${a#* };
  r=$(FI_B200_SCHED_F0=$1 FI_B200_SCHED_G=$2 FI_B200_SCHED_TAIL=$3 FI_B200_SCHED_LOOKAHEAD=$4 python bench.py --workload $w --shard-of $s --no-cpu --steps 200 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],2), round(d['roofline']['kernel_us_isolated'],1))")
  echo "F0=$1 G=$2 TAIL=$3 LOOK=$4 $w/$s: $r"; done; done
