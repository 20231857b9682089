#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 300 python tools/pcie_probe.py --gib 8 > $O/r2x_pcie_probe.json 2> $O/r2x_pcie_probe.err; echo "probe rc=$?"; cat $O/r2x_pcie_probe.json; tail -3 $O/r2x_pcie_probe.err
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "changing_buffers or repeat_bit" > $O/r2x_pytest.txt 2>&1; tail -3 $O/r2x_pytest.txt
nvidia-smi topo -m > $O/r2x_topo.txt 2>&1; lscpu | head -25 > $O/r2x_lscpu.txt; numactl -H >> $O/r2x_lscpu.txt 2>&1
nvidia-smi -q | grep -i -A6 "GPU Link Info\|PCIe Generation" | head -40 > $O/r2x_pcie_link.txt
cat $O/r2x_pcie_link.txt
