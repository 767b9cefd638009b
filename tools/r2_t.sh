python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | cut -c1-400
CMD="python tools/bench_qmv.py --bits 8 --M 32 --iters 3 --shapes gate_up,head"
$CMD > gpurun_out/qmi_plain.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:qmi -c 8 -o gpurun_out/prof_qmi $CMD > gpurun_out/ncu_qmi.log 2>&1
tail -2 gpurun_out/ncu_qmi.log
