"""Keep only the records of a tag range of an in-kernel trace dump (small enough to bring back from the GPU box)."""
import sys
import numpy as np
rec = np.fromfile(sys.argv[1], dtype=np.dtype([("t", "<u8"), ("tag", "<i4"), ("cta", "<i4")]))
lo, hi = int(sys.argv[3]), int(sys.argv[4])
print("records", len(rec), "tags", np.unique(rec["tag"]).tolist())
rec[(rec["tag"] >= lo) & (rec["tag"] < hi)].tofile(sys.argv[2])
