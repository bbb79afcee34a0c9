mkdir -p gpurun_out
timeout 300 python tools/prof_build.py C2 1.0 exact 0 1 2>&1 | tail -1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'sampler|flag_in|compact|degree' -c 40 --csv --log-file gpurun_out/r2_c2_sampler_launches.csv python tools/prof_build.py C2 1.0 exact 0 1 > /dev/null 2>&1
grep -E "sampler|flag_in|compact|degree" gpurun_out/r2_c2_sampler_launches.csv | awk -F'","' '{print $5, $(NF)}' | head -24
