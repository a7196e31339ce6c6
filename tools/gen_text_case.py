"""Write a seeded synthetic forest + data set as the reference's text files:
python tools/gen_text_case.py OUT_PREFIX TREES DEPTH COLS ROWS -> OUT_PREFIX.model.txt / OUT_PREFIX.data.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tahoe_b200 import synth
out, T, d, C, R = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
nodes = synth.gen_dense_forest(20240003, T, d, C)
data = synth.gen_data(20240103, R, C)
synth.write_model_text(out + ".model.txt", nodes, T, d)
synth.write_data_text(out + ".data.txt", data, -999.0)
print(f"wrote {out}: {T} trees depth {d}, {R} x {C}")
