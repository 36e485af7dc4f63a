"""Inception-v3 (arXiv:1512.00567) and Inception-v4 (arXiv:1602.07261).

Module tables follow T/models/inception_model.py:44-209: batch norm on every conv
(builder defaults: decay .999, eps 1e-3, no scale), factorised 1x7/7x1 and 1x3/3x1
convolutions, ``share`` columns, optional auxiliary head for v3.
"""
from . import model


class Inceptionv3Model(model.CNNModel):
    def __init__(self, auxiliary=False, params=None):
        self._auxiliary = auxiliary
        super().__init__("inception3", 299, 32, 0.005, params=params)

    def add_inference(self, cnn):
        def v3_a(n):
            cnn.inception_module("incept_v3_a", [
                [("conv", 64, 1, 1)],
                [("conv", 48, 1, 1), ("conv", 64, 5, 5)],
                [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3)],
                [("apool", 3, 3, 1, 1, "SAME"), ("conv", n, 1, 1)]])

        def v3_b():
            cnn.inception_module("incept_v3_b", [
                [("conv", 384, 3, 3, 2, 2, "VALID")],
                [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3, 2, 2, "VALID")],
                [("mpool", 3, 3, 2, 2, "VALID")]])

        def v3_c(n):
            cnn.inception_module("incept_v3_c", [
                [("conv", 192, 1, 1)],
                [("conv", n, 1, 1), ("conv", n, 1, 7), ("conv", 192, 7, 1)],
                [("conv", n, 1, 1), ("conv", n, 7, 1), ("conv", n, 1, 7), ("conv", n, 7, 1),
                 ("conv", 192, 1, 7)],
                [("apool", 3, 3, 1, 1, "SAME"), ("conv", 192, 1, 1)]])

        def v3_d():
            cnn.inception_module("incept_v3_d", [
                [("conv", 192, 1, 1), ("conv", 320, 3, 3, 2, 2, "VALID")],
                [("conv", 192, 1, 1), ("conv", 192, 1, 7), ("conv", 192, 7, 1),
                 ("conv", 192, 3, 3, 2, 2, "VALID")],
                [("mpool", 3, 3, 2, 2, "VALID")]])

        def v3_e(pooltype):
            cnn.inception_module("incept_v3_e", [
                [("conv", 320, 1, 1)],
                [("conv", 384, 1, 1), ("conv", 384, 1, 3)],
                [("share",), ("conv", 384, 3, 1)],
                [("conv", 448, 1, 1), ("conv", 384, 3, 3), ("conv", 384, 1, 3)],
                [("share",), ("share",), ("conv", 384, 3, 1)],
                [("mpool" if pooltype == "max" else "apool", 3, 3, 1, 1, "SAME"),
                 ("conv", 192, 1, 1)]])

        def v3_aux():
            assert cnn.aux_top_layer is None
            cnn.aux_top_layer, cnn.aux_top_size = cnn.top_layer, cnn.top_size
            with cnn.switch_to_aux_top_layer():
                cnn.apool(5, 5, 3, 3, mode="VALID")
                cnn.conv(128, 1, 1, mode="SAME")
                cnn.conv(768, 5, 5, mode="VALID", stddev=0.01)
                cnn.reshape([-1, 768])

        cnn.use_batch_norm = True
        cnn.conv(32, 3, 3, 2, 2, mode="VALID")   # 299 -> 149
        cnn.conv(32, 3, 3, 1, 1, mode="VALID")   # 147
        cnn.conv(64, 3, 3, 1, 1, mode="SAME")
        cnn.mpool(3, 3, 2, 2, mode="VALID")      # 73
        cnn.conv(80, 1, 1, 1, 1, mode="VALID")
        cnn.conv(192, 3, 3, 1, 1, mode="VALID")  # 71
        cnn.mpool(3, 3, 2, 2, "VALID")           # 35
        v3_a(32)
        v3_a(64)
        v3_a(64)
        v3_b()                                   # 17
        v3_c(128)
        v3_c(160)
        v3_c(160)
        v3_c(192)
        if self._auxiliary:
            v3_aux()
        v3_d()                                   # 8
        v3_e("avg")
        v3_e("max")
        cnn.apool(8, 8, 1, 1, "VALID")
        cnn.reshape([-1, 2048])


class Inceptionv4Model(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("inception4", 299, 32, 0.005, params=params)

    def add_inference(self, cnn):
        def sa():
            cnn.inception_module("incept_v4_sa", [
                [("mpool", 3, 3, 2, 2, "VALID")], [("conv", 96, 3, 3, 2, 2, "VALID")]])

        def sb():
            cnn.inception_module("incept_v4_sb", [
                [("conv", 64, 1, 1), ("conv", 96, 3, 3, 1, 1, "VALID")],
                [("conv", 64, 1, 1), ("conv", 64, 7, 1), ("conv", 64, 1, 7),
                 ("conv", 96, 3, 3, 1, 1, "VALID")]])

        def sc():
            cnn.inception_module("incept_v4_sc", [
                [("conv", 192, 3, 3, 2, 2, "VALID")], [("mpool", 3, 3, 2, 2, "VALID")]])

        def a():
            cnn.inception_module("incept_v4_a", [
                [("apool", 3, 3, 1, 1, "SAME"), ("conv", 96, 1, 1)],
                [("conv", 96, 1, 1)],
                [("conv", 64, 1, 1), ("conv", 96, 3, 3)],
                [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3)]])

        def ra(k, l, m, n):
            cnn.inception_module("incept_v4_ra", [
                [("mpool", 3, 3, 2, 2, "VALID")], [("conv", n, 3, 3, 2, 2, "VALID")],
                [("conv", k, 1, 1), ("conv", l, 3, 3), ("conv", m, 3, 3, 2, 2, "VALID")]])

        def b():
            cnn.inception_module("incept_v4_b", [
                [("apool", 3, 3, 1, 1, "SAME"), ("conv", 128, 1, 1)],
                [("conv", 384, 1, 1)],
                [("conv", 192, 1, 1), ("conv", 224, 1, 7), ("conv", 256, 7, 1)],
                [("conv", 192, 1, 1), ("conv", 192, 1, 7), ("conv", 224, 7, 1),
                 ("conv", 224, 1, 7), ("conv", 256, 7, 1)]])

        def rb():
            cnn.inception_module("incept_v4_rb", [
                [("mpool", 3, 3, 2, 2, "VALID")],
                [("conv", 192, 1, 1), ("conv", 192, 3, 3, 2, 2, "VALID")],
                [("conv", 256, 1, 1), ("conv", 256, 1, 7), ("conv", 320, 7, 1),
                 ("conv", 320, 3, 3, 2, 2, "VALID")]])

        def c():
            cnn.inception_module("incept_v4_c", [
                [("apool", 3, 3, 1, 1, "SAME"), ("conv", 256, 1, 1)],
                [("conv", 256, 1, 1)],
                [("conv", 384, 1, 1), ("conv", 256, 1, 3)],
                [("share",), ("conv", 256, 3, 1)],
                [("conv", 384, 1, 1), ("conv", 448, 1, 3), ("conv", 512, 3, 1),
                 ("conv", 256, 3, 1)],
                [("share",), ("share",), ("share",), ("conv", 256, 1, 3)]])

        cnn.use_batch_norm = True
        cnn.conv(32, 3, 3, 2, 2, mode="VALID")
        cnn.conv(32, 3, 3, 1, 1, mode="VALID")
        cnn.conv(64, 3, 3)
        sa()
        sb()
        sc()
        for _ in range(4):
            a()
        ra(192, 224, 256, 384)
        for _ in range(7):
            b()
        rb()
        for _ in range(3):
            c()
        cnn.spatial_mean()
        cnn.dropout(0.8)
