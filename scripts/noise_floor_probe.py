import sys, time, math
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from oracle.archs import ARCHS
from oracle import model as M
torch.set_num_threads(8)
a = ARCHS['llama-3.2-1b']
t=time.time(); m = M.OracleModel(a, 0, max_pos=512); print('built', time.time()-t, flush=True)
rng = np.random.default_rng(1); ids = rng.integers(0, a.vocab, 128).tolist()

def fwd(m, ids, ra, rq, perturb):
    # copy of forward with configurable rounding + accumulation perturbation
    a, sd = m.a, m.sd
    ids = torch.as_tensor(ids); T = ids.numel(); pos = torch.arange(T)
    x = sd['embed'][ids]; G = a.n_q//a.n_kv; scale = 1/math.sqrt(a.head_dim)
    def mm(h, w):
        if perturb:  # different summation order: split K in two halves
            K = h.shape[-1]//2
            return h[..., :K] @ w.T[:K] + h[..., K:] @ w.T[K:]
        return h @ w.T
    for l in range(a.layers):
        p=f'l{l}.'
        h, r = m._rms_split(x, sd[p+'in_norm']); h = ra(h)
        q = mm(h, sd[p+'q'])*r; k = mm(h, sd[p+'k'])*r; v = mm(h, sd[p+'v'])*r
        q = m._rope(q.view(T,a.n_q,a.head_dim), pos); k = m._rope(k.view(T,a.n_kv,a.head_dim), pos); v = v.view(T,a.n_kv,a.head_dim)
        k, v = M._rbf16(k), M._rbf16(v); q = rq(q)
        K_, V_ = k.transpose(0,1), v.transpose(0,1)
        qh = q.transpose(0,1).reshape(a.n_kv, G*T, a.head_dim)
        s = torch.bmm(qh, K_.transpose(1,2))*scale; s = s.view(a.n_kv,G,T,T)
        causal = torch.arange(T)[None,:] <= pos[:,None]
        s = s.masked_fill(~causal[None,None], float('-inf'))
        pr = torch.softmax(s, -1).view(a.n_kv, G*T, T)
        o = torch.bmm(pr, V_).view(a.n_kv,G,T,a.head_dim).permute(2,0,1,3).reshape(T, a.n_q*a.head_dim)
        x = x + mm(ra(o), sd[p+'o'])
        h, r = m._rms_split(x, sd[p+'post_norm']); h = ra(h)
        g = mm(h, sd[p+'gate'])*r; u = mm(h, sd[p+'up'])*r
        x = x + mm(ra(torch.nn.functional.silu(g)*u), sd[p+'down'])
    xf, r = m._rms_split(x, sd['final_norm'])
    return mm(ra(xf), sd['lm_head'])*r
bf = M._rbf16
f16 = lambda t: t.to(torch.float16).to(torch.float32)
idt = lambda t: t
for name, ra, rq in [('bf16 act, bf16 q', bf, bf), ('f16 act, bf16 q', f16, bf), ('f16 act, f16 q', f16, f16), ('exact act, bf16 q', idt, bf)]:
    t=time.time()
    A = fwd(m, ids, ra, rq, False); B = fwd(m, ids, ra, rq, True)
    rel = ((A-B).abs().amax(-1) / A.abs().amax(-1))
    rms = ((A-B).pow(2).mean(-1).sqrt() / A.abs().amax(-1))
    print(f'{name}: worst rel {rel.max():.3e} median {rel.median():.3e} rms {rms.max():.3e}  ({time.time()-t:.0f}s)', flush=True)
