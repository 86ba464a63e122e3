mkdir -p gpurun_out
for i in 1 2; do
AB_WORKLOADS=panoptic,droid timeout 300 python tools/ab.py mvtracker_b200/libmvt.so -- --no-model --no-distributions --no-adjacent --no-verify
echo "-- MVT_NO_DEFER=1 (plain kernel, nothing counted)"
MVT_NO_DEFER=1 AB_WORKLOADS=panoptic,droid timeout 300 python tools/ab.py mvtracker_b200/libmvt.so -- --no-model --no-distributions --no-adjacent --no-verify
done 2>&1 | tee gpurun_out/r2t_ab.txt
