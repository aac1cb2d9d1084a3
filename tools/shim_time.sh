cd /root/repo
python - <<PY
import sys
sys.path.insert(0,"tests"); sys.path.insert(0,".")
import bench, os
img=bench.make_corpus("config2",seed=2)
os.makedirs("/tmp/sh",exist_ok=True)
img.tofile("/tmp/sh/kernel-1.traceg"); open("/tmp/sh/kernelslist.g","w").write("kernel-1.traceg\n")
PY
cd /tmp/sh
for i in 1 2; do TI_SHIM_TIMING=1 /root/repo/oracle/_ref/tp_shim_cuda time ./kernelslist.g > /dev/null; done
