#!/bin/bash
# round 2, run 5: secondary kernels at A4 -- event times + ncu --set full counters -> profiles/r02_kernel_rooflines.json
set -u
mkdir -p gpurun_out
timeout 600 python tools/kernels_a4.py gpurun_out/r02n_kernel_events.json > /dev/null 2> gpurun_out/r02n_kernels.err; echo "plain rc=$?"; tail -2 gpurun_out/r02n_kernels.err
timeout 1500 ncu --set full --clock-control none -k regex:"gray_erode|resize_area|pack_bits|warp_flat|canny_|hough_acc|hough_peaks|fft_|jpeg_dct|jpeg_bits|jpeg_emit|jpeg_stuff_w|jdec_|sweep_finalize|sweeps_" -c 200 -f -o /tmp/r02n_kernels python tools/kernels_a4.py > gpurun_out/r02n_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02n_ncu.log
python tools/ncu_kernel_table.py gpurun_out/r02n_kernel_events.json /tmp/r02n_kernels.ncu-rep gpurun_out/r02n_kernel_rooflines.json
