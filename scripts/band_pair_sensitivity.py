"""CPU study behind the bars of tests/test_c5_gpu.py::test_stream_epilogues_and_fallback_streams_vs_scipy on the BANDED
synthetic pair: the reference recurrence itself (oracle.pcg / scripts/cg_order_sensitivity.cg) with nothing changed but
the summation order of its dot products moves by up to +3 iterations and its residual history by factors of 2-10
(non-monotone history with sharp dips), so there the test pins entries 0..3, the operator bits, the solution and a +-3
count; the unstructured pair is well behaved and keeps +-1 and the envelope."""
import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests'); sys.path.insert(0,'/root/repo/scripts')
from test_c5_gpu import _spd_pair
from test_pcg_gpu import history_within_envelope
from oracle import pcg as opcg
import cg_order_sensitivity as cs
for args in [(150000,4,7,True,True),(150000,4,7,False,True),(140000,4,3,True,True)]:
    A,G=_spd_pair(*args)
    rng=np.random.default_rng(1); rng.standard_normal(A.shape[0]); b=A@rng.standard_normal(A.shape[0])
    for kind in ("ext_spai","ext_spai_scaled"):
        M=opcg.make_preconditioner(kind,A,G,3e-3)
        xo,its,oh=opcg.pcg_restated(A,b,M,rtol=1e-8)
        def rev(a,c): return float(a[::-1]@c[::-1])
        def chunk(a,c):
            n=len(a); s=0.0
            for i in range(64): s+=float(a[i*n//64:(i+1)*n//64]@c[i*n//64:(i+1)*n//64])
            return s
        for nm,dot in (("rev",rev),("chunk",chunk)):
            out=cs.cg(A,b,M,dot,rtol=1e-8)
            h=out[-1] if isinstance(out,tuple) else out
            h=np.asarray(h); oh_=np.asarray(oh)
            m=min(len(h),len(oh_))
            rel=np.abs(h[:m]-oh_[:m])/oh_[:m]
            print(args,kind,nm,"its",len(h)-1,its,"max rel",rel.max(),"at",rel.argmax(),"env bad(2,.25)",history_within_envelope(h,oh_,2,0.25),"env(3,1.0)",history_within_envelope(h,oh_,3,1.0))
