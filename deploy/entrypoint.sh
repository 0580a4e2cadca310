#!/usr/bin/env bash
# Container entrypoint for both DDP topologies (/root/reference/README.md:21, 102, 120).
#   single-Pod Job:      torchrun --standalone --nproc_per_node=$NPROC_PER_NODE train.py ...
#   multi-Pod StatefulSet: NODE_RANK comes from the Pod ordinal (train-multipod-3 -> 3) and pod-0 is
#                          the rendezvous master, reached through the headless Service.
set -euo pipefail

NPROC_PER_NODE="${NPROC_PER_NODE:-1}"
NNODES="${NNODES:-1}"
MASTER_PORT="${MASTER_PORT:-29500}"

if [[ "${NNODES}" -gt 1 ]]; then
  host="${HOSTNAME:-$(hostname)}"
  NODE_RANK="${NODE_RANK:-${host##*-}}"
  if ! [[ "${NODE_RANK}" =~ ^[0-9]+$ ]]; then
    echo "entrypoint: cannot derive NODE_RANK from hostname '${host}'" >&2
    exit 2
  fi
  STS_NAME="${host%-*}"
  MASTER_ADDR="${MASTER_ADDR:-${STS_NAME}-0.${HEADLESS_SERVICE:-train-mp-headless}}"
  export NODE_RANK MASTER_ADDR MASTER_PORT
  # All Pods sit on ONE NVLink box: let NCCL treat them as one host so it uses NVLink P2P/NVLS
  # instead of the reference's TCP sockets (README.md:101).  Needs hostIPC + all GPUs visible.
  export NCCL_HOSTID="${NCCL_HOSTID:-nanosandbox-b200-box}"
  exec torchrun --nnodes="${NNODES}" --node_rank="${NODE_RANK}" --nproc_per_node="${NPROC_PER_NODE}" \
       --master_addr="${MASTER_ADDR}" --master_port="${MASTER_PORT}" "$@"
else
  exec torchrun --standalone --nproc_per_node="${NPROC_PER_NODE}" "$@"
fi
