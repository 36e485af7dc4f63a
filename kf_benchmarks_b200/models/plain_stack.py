"""Table-driven builder for the plain feed-forward nets of the zoo (LeNet, AlexNet, OverFeat,
the trivial FC model): a network is a list of layer records, each interpreted against the
NetBuilder. Records:

    ("conv", filters, kernel, stride, padding[, extra NetBuilder.conv keywords])
    ("maxpool", window, stride[, padding])
    ("lrn", radius, bias, alpha, beta)
    ("flatten",)            # [N, H, W, C] -> [N, H*W*C], size taken from the running shape
    ("fc", units[, extra NetBuilder.affine keywords])
    ("dropout",)
"""


def build_stack(cnn, plan):
    for rec in plan:
        kind, args = rec[0], rec[1:]
        if kind == "conv":
            filters, k, stride, padding = args[:4]
            extra = args[4] if len(args) > 4 else {}
            cnn.conv(filters, k, k, stride, stride, mode=padding, **extra)
        elif kind == "maxpool":
            window, stride = args[:2]
            cnn.mpool(window, window, stride, stride, mode=args[2] if len(args) > 2 else "VALID")
        elif kind == "lrn":
            cnn.lrn(*args)
        elif kind == "flatten":
            shape = cnn.top_layer.shape
            features = 1
            for d in shape[1:]:
                features *= d
            cnn.reshape([-1, features])
        elif kind == "fc":
            cnn.affine(args[0], **(args[1] if len(args) > 1 else {}))
        elif kind == "dropout":
            cnn.dropout()
        else:
            raise ValueError("unknown layer record %r" % (rec,))
