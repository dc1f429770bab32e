"""Layer ids of the vfi_debug_conv test hook (csrc/common.cuh layer table)."""
from .ifnet_params import BLOCK_C


def layer_info(layer):
    """-> dict(kind, cin, cout, cin_pad, cout_pad, stride, key) for layer id 0..58."""
    if layer < 4:
        tbl = [("conv", 3, 16, 4, 2, "encode.cnn0"), ("conv", 16, 16, 16, 1, "encode.cnn1"),
               ("conv", 16, 16, 16, 1, "encode.cnn2"), ("deconv", 16, 4, 16, 2, "encode.cnn3")]
        k, ci, co, cp, s, key = tbl[layer]
        return dict(kind=k, cin=ci, cout=co, cin_pad=cp, cout_pad=co, stride=s, key=key, beta=False, act=layer < 3)
    b, j = divmod(layer - 4, 11)
    c = BLOCK_C[b]
    p = f"block{b}."
    if j == 0:
        ci = 15 if b == 0 else 28
        return dict(kind="conv", cin=ci, cout=c // 2, cin_pad=16 if b == 0 else 32, cout_pad=c // 2, stride=2,
                    key=p + "conv0.0.0", beta=False, act=True)
    if j == 1:
        return dict(kind="conv", cin=c // 2, cout=c, cin_pad=c // 2, cout_pad=c, stride=2, key=p + "conv0.1.0",
                    beta=False, act=True)
    if j < 10:
        return dict(kind="conv", cin=c, cout=c, cin_pad=c, cout_pad=c, stride=1, key=p + f"convblock.{j - 2}.conv",
                    beta=True, act=True, beta_key=p + f"convblock.{j - 2}.beta")
    return dict(kind="deconv", cin=c, cout=52, cin_pad=c, cout_pad=52, stride=2, key=p + "lastconv.0", beta=False,
                act=False)


def layer_out_shape(layer, N, Hin, Win):
    li = layer_info(layer)
    if li["kind"] == "deconv":
        return (N, 2 * Hin, 2 * Win, 4) if layer == 3 else (N, 4 * Hin, 4 * Win, 16)
    s = li["stride"]
    return (N, Hin // s, Win // s, li["cout_pad"])
