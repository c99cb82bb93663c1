#!/bin/bash
# Probe of the GPU box for the reference's toolchain (VERDICT r1 "Next" 2a): Go + Lattigo module cache
{
echo "== date"; date -u
echo "== which go"; which go || echo "go: not found"
echo "== go version"; go version 2>&1 || true
echo "== go env GOMODCACHE"; go env GOMODCACHE 2>&1 || true
echo "== ls GOMODCACHE/github.com/tuneinsight"; ls "$(go env GOMODCACHE 2>/dev/null)/github.com/tuneinsight" 2>&1 || true
echo "== find lattigo"; find / -xdev -type d -iname 'lattigo*' 2>/dev/null | head
echo "== find go toolchains"; ls -d /usr/local/go /usr/lib/go* /root/go /opt/go* 2>&1 | head
echo "== javac / node"; which javac node 2>&1 || echo "none"
echo "== nproc"; nproc
echo "== nccl"; ls /usr/lib/x86_64-linux-gnu | grep -i nccl; python - <<'PY'
import torch, os
print("torch nccl", torch.cuda.nccl.version())
PY
nvidia-smi -L
} > gpurun_out/r2_go_probe.txt 2>&1
cat gpurun_out/r2_go_probe.txt
