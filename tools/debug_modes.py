"""Debug aid: the forward in e_mode 'store' vs 'scores' -- where do the two first differ?"""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'oracle'))
import torch
import pkgpath
import geotx_oracle as oracle

pkg = pkgpath.load_package()
DEV = 'cuda:0'
cfg = dict(pkg.synth.CONFIGS['3dmatch'])
rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair('3dmatch', 640, 512, seed=pkg.synth.SEED)
sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
model = pkg.GeometricTransformer(**cfg).to(DEV).eval()
model.load_state_dict(sd, strict=True)
args = [t.to(DEV) for t in (rp, sp, rf, sf)]
rec = {}
orig_att, orig_post = pkg.ops.attention, pkg.ops.post_attention


def rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm())


def run(mode):
    log = []
    def att(q, k, v, H, scores_p=None, key_mask=None, **kw):
        out = orig_att(q, k, v, H, scores_p, key_mask, **kw)
        log.append(('att', None if scores_p is None else scores_p.clone(), out.clone()))
        return out
    def post(hidden, x_in, *a, **kw):
        out = orig_post(hidden, x_in, *a, **kw)
        log.append(('post', hidden.clone(), out.clone()))
        return out
    pkg.ops.attention, pkg.ops.post_attention = att, post
    model.e_mode = mode
    with torch.no_grad():
        o = [t.clone() for t in model(*args)]
    pkg.ops.attention, pkg.ops.post_attention = orig_att, orig_post
    return o, log


o_s, log_s = run('store')
o_d, log_d = run('scores')
print('modes', model.last_e_mode, 'outputs', rel(o_d[0], o_s[0]), rel(o_d[1], o_s[1]))
for i, (a, b) in enumerate(zip(log_s, log_d)):
    sp_ = '' if a[1] is None or a[0] == 'post' else f'scores_p {rel(b[1], a[1]):.3e} (max abs {float((b[1]-a[1]).abs().max()):.3e}, |ref| max {float(a[1].abs().max()):.2f})'
    inp = f'hidden-in {rel(b[1], a[1]):.3e}' if a[0] == 'post' else sp_
    print(i, a[0], inp, 'out', f'{rel(b[2], a[2]):.3e}')
e0, e1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
print('vs oracle: store', rel(o_s[0].cpu(), e0), rel(o_s[1].cpu(), e1), 'scores', rel(o_d[0].cpu(), e0), rel(o_d[1].cpu(), e1))
