mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/prof_render.py c3 1080 64 dda 3 2>&1 | tail -2
timeout 300 python tools/prof_render.py c5 512 64 dda 3 2>&1 | tail -2
CMD="python tools/prof_render.py c3 1080 115 dda 1"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"k_wf_track|k_wf_nee" --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu_t.log 2>&1; echo "ncu traffic rc=$?"
python tools/ncu_traffic.py gpurun_out/traffic.csv gpurun_out/r1_tracking_traffic.json "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k regex:k_wf_track|k_wf_nee $CMD"
