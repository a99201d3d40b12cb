#!/usr/bin/env python
"""Node program of a decoder handle by (op, stage): tools/dump_ops.py KIND N K ctype L   (needs a GPU: the handle compiles the schedule)"""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import antpolarcodes_b200 as ap
kind, N, K, ct, L = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], int(sys.argv[5])
f = ap.frozen_bits(N, K, 0.0, ct)
c = ap._Code(kind, N, f, L, True, 16 if L > 1 else 0, 0)
ops = c.schedule()
names = {1: "F", 2: "G", 3: "C", 4: "G_R1", 5: "G_0R", 6: "C_0R", 7: "R0", 8: "R1", 9: "REP", 10: "REP2", 11: "SPC", 12: "0SPC", 13: "01", 14: "L_R0", 15: "L_R1", 16: "L_REP", 17: "L_SPC"}
cnt = collections.Counter((names.get(int(o) & 63, str(int(o) & 63)), (int(o) >> 6) & 15, (int(o) >> 10) & 7) for o in ops)
for k in sorted(cnt): print(k, cnt[k])
