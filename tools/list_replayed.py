#!/usr/bin/env python
"""Which scans of the synthetic VLP16 stream take the exact robin_map replay (lo_scan_stats.update_order)?
   python tools/list_replayed.py [scans]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 120
scans, _, _ = S.vlp16_stream(n, workers=8)
import lidar_odometry_demo_b200 as lo  # noqa: E402
ctx = lo.Context(0)
od = lo.LidarOdometry(ctx)
hits = []
for k, s in enumerate(scans):
    od.processCloud(s)
    if od.stats().update_order != lo.LO_ORDER_SORTED:
        hits.append((k, od.stats().update_order))
print("replayed / approximate scans:", hits)
