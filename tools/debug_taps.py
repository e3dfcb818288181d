import sys, torch, numpy as np
sys.path.insert(0, ".")
from oracle import convnet_oracle as O
from ddp_practice_b200 import Engine
from tests.refmodel import RefConvNet
from tests.helpers import grad_errors

dtype = sys.argv[1] if len(sys.argv) > 1 else "fp32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
amp = dtype != "fp32"
torch.manual_seed(3407)
st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
e = Engine(dtype=dtype, batch_cap=B, use_scaler=amp)
bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
e.load_state(st.flat_params().numpy(), bn.numpy())
x, y = O.synthetic_mnist(B * 4)
def rel(a, b): return ((a - b).norm() / b.norm()).item()
for s in range(4):
    xs, ys = x[s*B:(s+1)*B], y[s*B:(s+1)*B]
    # make the oracle start from the engine's exact parameters
    flat = e.params.cpu().clone(); off = 0
    for n, shp in O.PARAM_SPECS:
        k = int(np.prod(shp)); st.params[n] = flat[off:off+k].view(shp).clone(); off += k
    rp = O._RankPass(st, xs, ys, O._DT[dtype])
    m1 = rp.conv1(); m2 = rp.bn1_conv2(m1[0], m1[1]); rp.bn2_head(m2[0], m2[1])
    scale = st.scale if amp else 1.0
    s2 = rp.head_bwd(scale); s1 = rp.bn2_conv2_bwd(s2[0], s2[1], s2[0], s2[1], float(B*196))
    e.train_step(xs.cuda().contiguous(), ys.cuda().contiguous()); torch.cuda.synchronize()
    w2 = O.rnd(st.params["layer2.0.weight"], O._DT[dtype])
    dy2_ref = rp._bn_bwd_elemt(rp.da2, rp.y2, rp.mu2, rp.is2, st.params["layer2.1.weight"], s2[0], s2[1], float(B*196))
    dp1_ref = O.rnd(torch.nn.grad.conv2d_input(rp.p1.shape, w2, dy2_ref, 1, 2), O._DT[dtype])
    print(f"step {s}: y1 {rel(e.read_activation('y1',B), rp.y1):.2e} p1 {rel(e.read_activation('p1',B), rp.p1):.2e} "
          f"y2 {rel(e.read_activation('y2',B), rp.y2):.2e} p2 {rel(e.read_activation('p2',B), rp.p2):.2e} "
          f"dlogits {rel(e.read_activation('dlogits',B), rp.dlogits):.2e} dy2 {rel(e.read_activation('dy2',B), dy2_ref):.2e} "
          f"dp1 {rel(e.read_activation('dp1',B), dp1_ref):.2e}")
    d = (e.read_activation('dy2',B) - dy2_ref).abs()
    print("   dy2 max abs diff", d.max().item(), "at", np.unravel_index(d.argmax().item(), d.shape), "ref max", dy2_ref.abs().max().item(),
          "n elements > 1e-3*max:", int((d > 1e-3 * dy2_ref.abs().max()).sum()))
    a2diff = ((rp.a2 > 0) != (O._bn_apply(e.read_activation('y2',B), rp.mu2, rp.is2, st.params['layer2.1.weight'], st.params['layer2.1.bias'], O._DT[dtype]) > 0)).sum().item()
    print("   relu2 sign flips between engine-y2 and oracle-y2:", a2diff)
    st2 = st.clone()
    r = O.train_step(st, [xs], [ys], dtype=dtype, use_scaler=amp)
    if r.found_inf == 0:
        print("   grad errs", {k: f"{v:.1e}" for k, v in grad_errors(e.grads.cpu(), r.grads).items()})
