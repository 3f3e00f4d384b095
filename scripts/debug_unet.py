import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import hybridnet
from hybridnet import ops
from hybridnet.UNet import UNet
from oracle import nets
from tests.util import rel_l2, nchw_f32

sd = nets.init_state_dict("unet", 5, 0.2)
x = torch.rand(2, 3, 64, 64, generator=torch.Generator().manual_seed(1))
m = UNet(); m.load_state_dict(sd); m = m.cuda().train()
ctx = nets.Ctx(training=True)
xn = ops.to_nhwc(x.cuda())
with torch.no_grad():
    # oracle stepwise
    e1 = nets._unet_block(sd, "encoder1", "enc1", x, ctx)
    p1 = F.max_pool2d(e1, 2)
    e2 = nets._unet_block(sd, "encoder2", "enc2", p1, ctx)
    e3 = nets._unet_block(sd, "encoder3", "enc3", F.max_pool2d(e2, 2), ctx)
    e4 = nets._unet_block(sd, "encoder4", "enc4", F.max_pool2d(e3, 2), ctx)
    bt = nets._unet_block(sd, "bottleneck", "bottleneck", F.max_pool2d(e4, 2), ctx)
    up4 = F.conv_transpose2d(bt, sd["upconv4.weight"], sd["upconv4.bias"], stride=2)
    d4 = nets._unet_block(sd, "decoder4", "dec4", torch.cat((up4, e4), 1), ctx)
    rb = UNet._run_block
    E1 = rb(m.encoder1, xn); print("enc1", rel_l2(nchw_f32(E1, 32).cpu(), e1))
    # first conv only
    z = F.conv2d(x, sd["encoder1.enc1conv1.weight"], None, 1, 1)
    Z = ops.conv2d(xn, m.encoder1[0].weight, None)
    print("conv1 raw", rel_l2(nchw_f32(Z, 32).cpu(), z))
    Zs, s0, s1 = ops._ConvStatsFn.apply(xn, None, m.encoder1[0].weight, None, 1, 1, 1, 0, False)
    print("conv1 stats-variant", rel_l2(nchw_f32(Zs, 32).cpu(), z))
    P = z.numel() / 32
    print("mean err", (s0[:32].cpu() / P - z.mean(dim=(0, 2, 3))).abs().max().item(), "sq err",
          (s1[:32].cpu() / P - (z * z).mean(dim=(0, 2, 3))).abs().max().item())
    E2 = rb(m.encoder2, ops.maxpool2(E1)); print("enc2", rel_l2(nchw_f32(E2, 64).cpu(), e2))
    E3 = rb(m.encoder3, ops.maxpool2(E2)); print("enc3", rel_l2(nchw_f32(E3, 128).cpu(), e3))
    E4 = rb(m.encoder4, ops.maxpool2(E3)); print("enc4", rel_l2(nchw_f32(E4, 256).cpu(), e4))
    BT = rb(m.bottleneck, ops.maxpool2(E4)); print("bott", rel_l2(nchw_f32(BT, 512).cpu(), bt))
    U4 = m.upconv4.nhwc(BT); print("up4", rel_l2(nchw_f32(U4, 256).cpu(), up4))
    D4 = rb(m.decoder4, U4, x2=E4); print("dec4", rel_l2(nchw_f32(D4, 256).cpu(), d4))
