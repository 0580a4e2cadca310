#!/usr/bin/env bash
# BASELINE.json configs[4]: GPT-2 medium as an 8-Pod StatefulSet (nnodes=8 x nproc_per_node=1) on ONE NVLink box
# (/root/reference/README.md:8, 62-72, 101-102).  There is no kubelet here, so the eight "Pods" are eight
# separate launches of the container entrypoint (deploy/entrypoint.sh), each with the environment the
# StatefulSet gives its Pod: HOSTNAME=train-multipod-<ordinal> (-> NODE_RANK), one visible GPU, the headless
# service name replaced by 127.0.0.1.  Every Pod runs its own `torchrun --nnodes=8 --node_rank=i
# --nproc_per_node=1`, exactly as k8s/statefulset/40-train-multipod.yaml does.
#
#   bash tools/multipod_emulation.sh [out_dir] [-- extra bench.py args]
#
# Writes <out_dir>/multipod_bench.json (rank 0's bench line), pod_<i>.log and NCCL's own account of the
# transports it chose (NCCL_DEBUG=INFO): the question the reference's TCP-socket setup raises is whether NCCL
# reaches the other Pods' GPUs over NVLink (P2P / NVLS) or falls back to sockets.
set -uo pipefail
OUT="${1:-gpurun_out/multipod}"
shift || true
[[ "${1:-}" == "--" ]] && shift
mkdir -p "$OUT"
NPODS="${NPODS:-8}"
PORT="${MASTER_PORT:-29517}"
pids=()
for i in $(seq 0 $((NPODS - 1))); do
  (
    export CUDA_VISIBLE_DEVICES="$i" HOSTNAME="train-multipod-$i" NNODES="$NPODS" NPROC_PER_NODE=1
    export MASTER_ADDR=127.0.0.1 MASTER_PORT="$PORT"
    export NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,GRAPH,TUNING NCCL_DEBUG_FILE="$OUT/nccl_pod$i.%p.log"
    if [[ "$i" == 0 ]]; then
      bash deploy/entrypoint.sh bench.py --config gpt2-medium --gpus "$NPODS" --steps 5 --warmup 3 \
           --no-cpu-baseline --no-torch-baseline "$@" > "$OUT/multipod_bench.json" 2> "$OUT/pod_$i.log"
    else
      bash deploy/entrypoint.sh bench.py --config gpt2-medium --gpus "$NPODS" --steps 5 --warmup 3 \
           --no-cpu-baseline --no-torch-baseline "$@" > "$OUT/pod_$i.out" 2> "$OUT/pod_$i.log"
    fi
  ) &
  pids+=($!)
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=$?; done
echo "pods finished rc=$rc"
# what NCCL says it is using between the Pods
{
  echo "# NCCL transports between the $NPODS emulated Pods (NCCL_DEBUG=INFO, all ranks)"
  cat "$OUT"/nccl_pod*.log 2>/dev/null | grep -E "NCCL version|via P2P|via SHM|via NET|NVLS|Connected all|Channel [0-9]+/[0-9]+ :|nNodes|comm 0x.* rank .* nranks" \
    | sed -E 's/^[^ ]+ //' | sort | uniq -c | sort -rn | head -60
} > "$OUT/nccl_transport_summary.txt"
cat "$OUT/nccl_transport_summary.txt" | head -40
cat "$OUT/multipod_bench.json"
exit $rc
