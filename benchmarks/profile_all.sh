#!/bin/bash
# ncu captures of the remaining fused paths (run on ONE GPU; the plain run must exit 0 first).
set -e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python benchmarks/profile_kernels.py > gpurun_out/profile_plain.log 2>&1
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:grad_dense_rowwarp_kernel -c 1 -o gpurun_out/prof_rowwarp python benchmarks/profile_kernels.py --which rowwarp > gpurun_out/ncu_rowwarp.log 2>&1
$NCU -k regex:grad_dense_cluster_kernel -c 1 -o gpurun_out/prof_cluster python benchmarks/profile_kernels.py --which cluster > gpurun_out/ncu_cluster.log 2>&1
$NCU -k regex:grad_dense_kernel -c 1 -o gpurun_out/prof_bf16 python benchmarks/profile_kernels.py --which bf16 > gpurun_out/ncu_bf16.log 2>&1
$NCU -k regex:grad_ell_smem_kernel -c 1 -o gpurun_out/prof_ell_smem python benchmarks/profile_kernels.py --which ell_smem > gpurun_out/ncu_ell_smem.log 2>&1
$NCU -k regex:grad_ell_kernel -c 1 -o gpurun_out/prof_ell_global python benchmarks/profile_kernels.py --which ell_global > gpurun_out/ncu_ell_global.log 2>&1
timeout 300 $NCU -k regex:decoder_round_kernel -s 3 -c 1 -o gpurun_out/prof_decoder python benchmarks/profile_kernels.py --which round > gpurun_out/ncu_decoder.log 2>&1 || echo decoder capture failed
tail -2 gpurun_out/ncu_*.log
