"""SSD-300 with the MLPerf ResNet-34 backbone (``--model=ssd300 --data_name=coco``).

Reference: T/models/ssd_model.py:47-552, constants T/ssd_constants.py.

* backbone = ResNet-34 v1 without its 4th group and with stride 1 in the 3rd
  (:96-136): 7x7/2 conv, 3x3/2 max-pool, basic blocks 3x64, 4x128 (first strided, projection
  shortcut), 6x256 (projection shortcut on the first) -> 38x38x256; BN decay .997, eps 1e-4;
* five extra 1x1 -> 3x3 conv pairs with bias+ReLU give maps 19, 10, 5, 3, 1 (:152-171);
* each of the six maps gets a 3x3 location head (4 x D channels) and a 3x3 class head
  (81 x D channels), D = 4,6,6,6,4,4 default boxes -> 8732 anchors (:173-193);
* loss = smooth-L1 on matched anchors + cross-entropy with 3:1 hard-negative mining
  (:299-384), computed by ``csrc/loss_kernels.cu`` straight from the twelve head outputs
  (the reference's [N, 8732, 85] concat / transpose is never materialised);
* learning rate 1e-3 at batch 32, linear in the batch size, /num_gpus in replicated mode,
  5 warm-up epochs, x0.1 at 160k and 200k steps of batch 32 (:223-256).

Synthetic inputs follow :541-552 (uniform images, boxes and classes in [0,1), 1..9 matched
boxes). With ``--data_dir`` the COCO TFRecords are read, augmented and labelled by
``data/ssd_dataloader.py`` (NumPy restatement of T/ssd_dataloader.py and of the pieces of the
external object_detection package it uses); ``--eval`` decodes the heads, runs per-class NMS and
computes COCO AP/AR with ``coco_metric.py`` (NumPy restatement of the COCOeval protocol -
pycocotools is not available offline). ``--backbone_model_path`` loads the ResNet-34 backbone
from a checkpoint of this framework or from an npz keyed like the TF official ResNet (there is
no reader for TF's binary checkpoint format: convert with ``np.savez`` first).
"""
import torch

from . import model, resnet_model
from ..nn.builder import NetBuilder
from ..nn.graph import Graph

NUM_CLASSES = 81
NUM_SSD_BOXES = 8732
NUM_DEFAULTS = (4, 6, 6, 6, 4, 4)
FEATURE_SIZES = (38, 19, 10, 5, 3, 1)
NEGS_PER_POSITIVE = 3
BATCH_NORM_DECAY = 0.997
BATCH_NORM_EPSILON = 1e-4
COCO_NUM_TRAIN_IMAGES = 118287


class _SsdSyntheticSource:
    def __init__(self, batch, image_size, device, seed, host=False):
        gen = torch.Generator().manual_seed(int(seed))
        self.tensors = [
            torch.rand((batch, image_size, image_size, 3), generator=gen).half(),
            torch.rand((batch, NUM_SSD_BOXES, 4), generator=gen),
            torch.rand((batch, NUM_SSD_BOXES, 1), generator=gen),
            torch.rand((batch,), generator=gen) * 9.0 + 1.0,
        ]
        dev = torch.device(device)
        if dev.type == "cuda" and not host:
            self.tensors = [t.to(dev) for t in self.tensors]
        elif dev.type == "cuda":
            self.tensors = [t.pin_memory() for t in self.tensors]
            self.host_resident = True
        self.bytes_per_step = sum(t.numel() * t.element_size() for t in self.tensors)

    def next_batch(self):
        return tuple(self.tensors)


class SSD300Model(model.CNNModel):
    def __init__(self, label_num=NUM_CLASSES, batch_size=32, learning_rate=1e-3,
                 backbone="resnet34", params=None):
        super().__init__("ssd300", 300, batch_size, learning_rate, params=params)
        if backbone != "resnet34":
            raise ValueError("Invalid backbone model %s for SSD." % backbone)
        self.label_num = label_num
        self.out_chan = [256, 512, 512, 256, 256, 256]
        self.num_dboxes = list(NUM_DEFAULTS)
        self.base_lr_batch_size = 32
        self.aux_loss_weight = 0.0

    def skip_final_affine_layer(self):
        return True

    def add_backbone_model(self, cnn):
        cnn.conv(64, 7, 7, 2, 2, mode="SAME_RESNET", use_batch_norm=True)
        cnn.mpool(3, 3, 2, 2, mode="SAME")
        for _ in range(3):
            resnet_model.residual_block(cnn, 64, 1, "v1")
        for i in range(4):
            resnet_model.residual_block(cnn, 128, 2 if i == 0 else 1, "v1", i == 0)
        for i in range(6):
            resnet_model.residual_block(cnn, 256, 1, "v1", i == 0)

    def add_inference(self, cnn):
        cnn.use_batch_norm = True
        cnn.batch_norm_config = {"decay": BATCH_NORM_DECAY, "epsilon": BATCH_NORM_EPSILON,
                                 "scale": True}
        with cnn.scope("resnet34_backbone"):
            self.add_backbone_model(cnn)

        def ssd_layer(depth, k, stride, mode):
            return cnn.conv(depth, k, k, stride, stride, mode=mode, use_batch_norm=False)
        activations = [cnn.top_layer]
        ssd_layer(256, 1, 1, "valid")
        activations.append(ssd_layer(512, 3, 2, "same"))
        ssd_layer(256, 1, 1, "valid")
        activations.append(ssd_layer(512, 3, 2, "same"))
        ssd_layer(128, 1, 1, "valid")
        activations.append(ssd_layer(256, 3, 2, "same"))
        ssd_layer(128, 1, 1, "valid")
        activations.append(ssd_layer(256, 3, 1, "valid"))
        ssd_layer(128, 1, 1, "valid")
        activations.append(ssd_layer(256, 3, 1, "valid"))
        self.loc, self.conf = [], []
        for nd, ac, oc in zip(self.num_dboxes, activations, self.out_chan):
            self.loc.append(cnn.conv(nd * 4, 3, 3, 1, 1, input_layer=ac, num_channels_in=oc,
                                     activation=None, use_batch_norm=False))
            self.conf.append(cnn.conv(nd * self.label_num, 3, 3, 1, 1, input_layer=ac,
                                      num_channels_in=oc, activation=None,
                                      use_batch_norm=False))
        cnn.top_size = 4 + self.label_num

    def build_graph(self, nclass=None, phase_train=True, batch_size=None) -> Graph:
        n = batch_size or self.batch_size
        g = Graph(n, (self.image_size, self.image_size, self.depth))
        cnn = NetBuilder(g, phase_train=phase_train)
        self.add_inference(cnn)
        sizes = [t.shape[1] for t in self.loc]
        anchors = sum(s * s * d for s, d in zip(sizes, self.num_dboxes))
        g.label_specs = [((n, anchors, 4), "f32"), ((n, anchors, 1), "f32"), ((n,), "f32")]
        g.loss = {"kind": "ssd", "loc": [t.id for t in self.loc],
                  "conf": [t.id for t in self.conf], "num_dboxes": list(self.num_dboxes),
                  "num_classes": self.label_num, "negs_per_positive": NEGS_PER_POSITIVE}
        g.logits = None
        return g

    def get_input_shapes(self, subset="train"):
        del subset
        return [[self.batch_size, self.image_size, self.image_size, self.depth],
                [self.batch_size, NUM_SSD_BOXES, 4], [self.batch_size, NUM_SSD_BOXES, 1],
                [self.batch_size]]

    def get_scaled_base_learning_rate(self, batch_size):
        base_lr = self.learning_rate
        if self.params is not None and self.params.variable_update == "replicated":
            base_lr = self.learning_rate / self.params.num_gpus
        return base_lr * (batch_size / self.base_lr_batch_size)

    def l2_decay_filter(self, var_name: str) -> bool:
        """T/models/ssd_model.py:91-94: no variable whose name contains 'batchnorm' is decayed."""
        return "batchnorm" not in var_name

    def get_learning_rate(self, global_step, batch_size):
        lr0 = self.get_scaled_base_learning_rate(batch_size)
        boundaries = [b * self.base_lr_batch_size // batch_size for b in (160000, 200000)]
        warmup_steps = int(COCO_NUM_TRAIN_IMAGES / batch_size * 5)
        if global_step < warmup_steps:
            return lr0 * float(global_step) / float(warmup_steps)
        sched = model.LRSchedule([lr0, lr0 * 0.1, lr0 * 0.01], boundaries)
        return sched(global_step)

    def make_input_source(self, batch, device, seed, rank=0, host=False, world=1, train=True):
        if self.params is not None and self.params.data_dir:
            from ..data.ssd_dataloader import CocoInputSource
            return CocoInputSource(self.params.data_dir, batch, device, seed, rank, world, train)
        return _SsdSyntheticSource(batch, self.image_size, device, seed * 1000003 + rank,
                                   host=host)

    def load_backbone_model(self, store, path) -> int:
        """Initialises the ResNet-34 backbone from a checkpoint (T/models/ssd_model.py:256-297,
        386-394). Accepted layouts: a checkpoint of this framework (variables named
        ``v0/cg/resnet34_backbone/...``) or an npz keyed like the TF official ResNet the
        reference expects (``conv2d[_N]/kernel``, ``batch_normalization[_N]/gamma`` ...), matched
        by layer index exactly as ``_var_name_in_official_model_ckpt`` does. Returns the number
        of variables loaded."""
        import re
        from .. import ckpt as ckpt_lib
        arrays = ckpt_lib.load_arrays(path)
        wanted = {}
        for name in store.graph.params:
            if "resnet34_backbone" not in name:
                continue
            key = "v0/" + name
            if key in arrays:
                wanted[key] = arrays[key]
                continue
            m = re.search(r"conv(\d+)/conv2d/(.+)$", name)
            official = None
            if m:
                official = "conv2d" + ("_%s" % m.group(1) if int(m.group(1)) else "") + "/" + \
                    m.group(2)
            else:
                m = re.search(r"batchnorm(\d+)/(.+)$", name)
                if m:
                    official = "batch_normalization" + \
                        ("_%s" % m.group(1) if int(m.group(1)) else "") + "/" + m.group(2)
            for cand in (official, "resnet_model/%s" % official if official else None):
                if cand and cand in arrays:
                    wanted[key] = arrays[cand]
                    break
        if not wanted:
            raise ValueError("no backbone variables found in %s" % path)
        store.from_tf_arrays(wanted, strict=False)
        return len(wanted)

    # -- evaluation: decode the heads, collect detections, COCO mAP at the end -------------
    def predict(self, executor):
        """(boxes [N, 8732, 4] decoded + normalised, scores [N, 8732, 81]) of the last forward
        pass (T/models/ssd_model.py:455-479)."""
        from ..data import ssd_dataloader as sd
        from ..nn.losses_ref import ssd_gather
        g = executor.graph
        vals = {tid: executor.value(tid) for tid in list(g.loss["loc"]) + list(g.loss["conf"])}
        loc, conf = ssd_gather(g, vals)
        scores = torch.softmax(conf.float(), dim=2).cpu().numpy()
        boxes = sd.decode_boxes(loc.float().cpu().numpy(), sd.default_boxes()("ltrb"))
        return boxes, scores

    def eval_step(self, executor, batch, src):
        executor.set_primary_input(batch[0])
        executor.forward(False)
        boxes, scores = self.predict(executor)
        meta = getattr(src, "last_eval_meta", None) or [(i, (300, 300, 3))
                                                        for i in range(len(boxes))]
        if not hasattr(self, "predictions"):
            self.predictions = {}
        for i, (sid, raw) in enumerate(meta):
            self.predictions[int(sid)] = {"pred_boxes": boxes[i], "pred_scores": scores[i],
                                          "source_id": int(sid), "raw_shape": raw}
        return 0.0

    def eval_finish(self, params):
        """COCO mAP over everything collected (T/models/ssd_model.py:481-539)."""
        import os
        from .. import coco_metric
        preds = list(getattr(self, "predictions", {}).values())
        self.predictions = {}
        ann = os.path.join(params.data_dir or "", "annotations", "instances_val2017.json")
        if not preds or not os.path.exists(ann):
            return {"top_1_accuracy": 0.0, "top_5_accuracy": 0.0,
                    "detections_collected": len(preds)}
        from .. import constants
        # like the reference (T/models/ssd_model.py:531-535) the metrics are marked as simple
        # values, which sends them to the eval summary as eval/COCO/AP ...
        out = {constants.SIMPLE_VALUE_RESULT_PREFIX + k: v
               for k, v in coco_metric.compute_map(preds, ann).items()}
        out.update({"top_1_accuracy": 0.0, "top_5_accuracy": 0.0})
        return out

    def postprocess(self, results):
        return results
