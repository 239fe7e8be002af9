#!/bin/sh
# One node, one rank per B200 (the reference's start_ddp.sh starts node 0 of 4 single-process hosts on 172.18.0.2:6585).
# Override NPROC / NNODES / NODE_RANK / MASTER_ADDR to reproduce that multi-host form.
exec torchrun --nproc_per_node="${NPROC:-8}" --nnodes="${NNODES:-1}" --node_rank="${NODE_RANK:-0}" \
  --master_addr="${MASTER_ADDR:-127.0.0.1}" --master_port="${MASTER_PORT:-6585}" main_ddp.py "$@"
