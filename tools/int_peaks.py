import ctypes as C, json, sys
sys.path.insert(0,'.')
from ydorbslam_b200._lib import lib, check
out=(C.c_double*7)()
check(lib().ydorb_diag_int_peaks(0,out,7))
names=["popc","lop3","vimnmx3_s16x2","imad","lop3_imad_mix","hmnmx2","hmnmx2_vimnmx3_mix"]
d={n:out[i]/1e9 for i,n in enumerate(names)}
d["unit"]="G thread-ops/s"
d["per_clk_per_sm_at_1965MHz"]={n:out[i]/(148*1.965e9) for i,n in enumerate(names)}
print(json.dumps(d))
