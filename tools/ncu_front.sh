for k in dds_sep_ln_gelu_kernel layernorm_kernel attention96_kernel conv_f32_w192_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o gpurun_out/fe_$k -f python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_fe_$k.log 2>&1
  ncu -i gpurun_out/fe_$k.ncu-rep --page details --csv > gpurun_out/fe_$k.details.csv 2>/dev/null
done
ls -la gpurun_out/*.ncu-rep
