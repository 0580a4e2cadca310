#!/usr/bin/env bash
# One 8-GPU box: the headline workload at N=8 (with A/B of the tile-scheduler policy and of the NCCL CTA cap),
# the char workload at N=8 (BASELINE.json configs[2]) and the 8-Pod emulation of configs[4].
#   bash tools/scale_run.sh [out_dir]
OUT="${1:-gpurun_out/scale}"
mkdir -p "$OUT"
tr() { # name, extra env..., -- bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 200)) bench.py --gpus 8 --no-cpu-baseline "$@" > "$OUT/$name.json" 2> "$OUT/$name.err"
  echo "$name rc=$?"; cut -c1-400 "$OUT/$name.json"
  python - "$OUT/$name.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    a = d.get("allreduce") or {}
    print("   value", round(d["value"]), "ms/step", round(d["ms_per_step"], 2), "gemm avg us", round(d["roofline"]["avg_launch_us"], 2),
          "frac", round(d["roofline"]["frac"], 3), "exposed_ms", a.get("exposed_comm_ms"), "ddp_overhead_ms", a.get("ddp_overhead_ms"),
          "comm_stream_ms", a.get("ms_on_comm_stream"), "ddp_check", (d.get("ddp_check") or {}).get("ok"))
except Exception as e:
    print("   (no result line:", e, ")")
PY
}
tr n8_default NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING -- --steps 20 --warmup 5
tr n8_static_tiles NSB_GEMM_DYNAMIC=0 -- --steps 20 --warmup 5
tr n8_no_cta_cap NSB_NCCL_MAX_CTAS=0 -- --steps 20 --warmup 5
tr n8_char -- --config char --steps 200 --warmup 20
grep -E "NCCL version|ncclCommSplit|NVLS multicast|Connected all|Connected NVLS|via P2P" "$OUT/n8_default.err" | sed -E 's/^[^ ]+ \[[0-9]\] //' | sed -E 's/comm 0x[0-9a-f]+/comm/; s/commId 0x[0-9a-f]+//; s/busId [0-9a-f]+//; s/parent 0x[0-9a-f]+//' | sort | uniq -c | sort -rn | head -20 > "$OUT/n8_nccl_summary.txt"
bash tools/multipod_emulation.sh "$OUT/multipod" > "$OUT/multipod_run.log" 2>&1; tail -45 "$OUT/multipod_run.log" | cut -c1-400
