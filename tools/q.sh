mkdir -p gpurun_out/mpc
O=gpurun_out/mpc
for P in 256 1024; do
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/traffic_nc_$P.csv python tools/mpc_traffic.py $P > $O/ncu_nc_$P.log 2>&1; echo rc=$?
done
