#!/bin/bash
# round 2, GPU call 14 (2 GPUs): ncu on the single-process two-device commit -- the fused scatter's peer traffic (NVLink) on the last NTT pass
mkdir -p gpurun_out
ncu --query-metrics 2>/dev/null | grep -i -E "nvl|peer" | head -60 > gpurun_out/r02_ncu_peer_metric_names.txt
wc -l gpurun_out/r02_ncu_peer_metric_names.txt
python tools/multi_probe.py 20 64 > gpurun_out/plain_multi.log 2>&1 && \
ncu --target-processes all --clock-control none -k regex:k_ntt_pass -c 24 --metrics gpu__time_duration.sum,lts__t_sectors_srcunit_tex_aperture_peer_op_write.sum,lts__t_sectors_aperture_peer_op_write.sum,lts__t_bytes_aperture_peer_op_write.sum,l1tex__m_xbar2l1tex_read_bytes.sum,dram__bytes_write.sum,nvltx__bytes.sum,nvlrx__bytes.sum --csv --log-file gpurun_out/r02_ncu_scatter_n2.csv python tools/multi_probe.py 20 64 > gpurun_out/ncu_multi.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_multi.log; head -c 1500 gpurun_out/r02_ncu_scatter_n2.csv
