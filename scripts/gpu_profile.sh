#!/bin/bash
# ncu captures of the headline kernels at their benchmark shapes (one GPU); the raw CSV comes back in
# gpurun_out/ and is summarised into profiles/ (see profiles/README.md).
mkdir -p gpurun_out
timeout 600 python scripts/profile_kernels.py all > gpurun_out/plain_prof.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none -k regex:'sv_forward_kernel|sv_adjoint_kernel|unitary_gemm_kernel|fedavg_fused_kernel|sv_large_pass_kernel|sv_global_swap_fused_kernel|pack_product' -c 30 -o /tmp/prof_kernels python scripts/profile_kernels.py all > gpurun_out/ncu_kernels.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_kernels.log | cut -c1-200
ncu -i /tmp/prof_kernels.ncu-rep --page raw --csv > gpurun_out/prof_kernels_raw.csv 2>/dev/null; ls -la /tmp/prof_kernels.ncu-rep gpurun_out/prof_kernels_raw.csv
