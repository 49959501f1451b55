"""Times the config-4 Hamming top-2 (2000 x 1M) for a few grid sizes (YDORB_TOP_CTAS_PER_SM knob)."""
import os
import subprocess
import sys

if len(sys.argv) > 1 and sys.argv[1] == "one":
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from ydorbslam_b200 import OrbMatcher, synth
    dev = torch.device("cuda:0")
    m = OrbMatcher(0.6, device=0)
    st = torch.cuda.Stream()
    m.set_stream(st.cuda_stream)
    nq, ndb = 2000, 1_000_000
    d_db = torch.from_numpy(synth.synth_descriptors(ndb, 1000)).to(dev)
    d_q = torch.from_numpy(synth.synth_descriptors(nq, 7)).to(dev)
    out = torch.zeros(nq, 4, dtype=torch.int32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for _ in range(3):
            m.top2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, 0, out.data_ptr())
        e0.record(st)
        for _ in range(10):
            m.top2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, 0, out.data_ptr())
        e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(os.environ.get("YDORB_TOP_CTAS_PER_SM"), f"{ms:.3f} ms  {nq * ndb / ms / 1e6:.1f} Gpairs/s")
else:
    for c in sys.argv[1:] or ["2", "4", "6", "8", "12"]:
        subprocess.run([sys.executable, __file__, "one"], env=dict(os.environ, YDORB_TOP_CTAS_PER_SM=c))
