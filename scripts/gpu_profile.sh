#!/bin/bash
# ncu captures of the headline kernels (one GPU; never under a multi-rank launcher).  The plain run must exit 0 first.
# The report itself stays on the box (well over the 64 MiB that come back); its raw page and the source page of the
# persistent TinyCNN kernel come back as CSV in gpurun_out/; scripts/summarise_ncu.py condenses the raw page into profiles/.
mkdir -p gpurun_out
WHICH=${1:-all}
KREGEX='sv_forward_kernel|sv_adjoint_kernel|unitary_gemm_kernel|fedavg_fused_kernel|fedavg_fused_loopback_kernel|sv_large_pass|sv_large_adjoint|sv_large_parity|sv_large_expz|sv_global_swap_fused_kernel|pack_product|tinycnn_round_kernel|tinycnn_eval_conv_kernel|head_eval|vqc_local_train_kernel'
timeout 600 python scripts/profile_kernels.py $WHICH > gpurun_out/plain_prof.log 2>&1 && \
timeout 2400 ncu --set full --clock-control none --import-source on -k regex:"$KREGEX" -c 60 -o /tmp/prof_kernels python scripts/profile_kernels.py $WHICH > gpurun_out/ncu_kernels.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_kernels.log | cut -c1-200
ncu -i /tmp/prof_kernels.ncu-rep --page raw --csv > gpurun_out/prof_kernels_raw.csv 2>/dev/null
ncu -i /tmp/prof_kernels.ncu-rep --page source --csv -k regex:tinycnn_round_kernel 2>/dev/null | head -c 20000000 > gpurun_out/prof_tinycnn_round_source.csv
ls -la /tmp/prof_kernels.ncu-rep gpurun_out/
