"""gather-bandwidth probe over buffer sizes: where the L2 -> SM rate ends and the HBM / TLB-bound rate begins"""
import os
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402

vx = bench.pkg("vx")
dev = vx.Device(0)
for mb in (4, 16, 32, 64, 96, 128, 256, 512, 1024, 2048):
    print(f"{mb:5d} MB   32 B records {dev.probe_gather(mb << 20, 32):8.0f} GB/s   128 B records {dev.probe_gather(mb << 20, 128):8.0f} GB/s")
dev.close()
