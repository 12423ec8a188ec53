O=gpurun_out/last
mkdir -p $O
A="heat3d --depth 128 --height 512 --width 1024 --generations 1 --cluster3d"
timeout 40 python scripts/ncu_case.py $A > $O/plain_cluster.log 2>&1 && timeout 60 ncu --set full --clock-control none -k regex:step3d_cluster -s 1 -c 1 -o $O/prof_heat3d_cluster -f python scripts/ncu_case.py $A > $O/ncu_cluster.log 2>&1
python scripts/ncu_summary.py $O/prof_heat3d_cluster.ncu-rep > $O/prof_heat3d_cluster.summary.txt 2>&1; rm -f $O/*.ncu-rep
