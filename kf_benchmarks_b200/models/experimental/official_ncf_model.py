"""NCF / NeuMF recommendation model (``--model=ncf``).

Reference: T/models/experimental/official_ncf_model.py:36-129 wraps
``official.recommendation.neumf_model.construct_model`` of the external tensorflow/models
repository with the MovieLens-20M sizes (138 493 users x 26 744 items), MLP layers
(256, 256, 128, 64), mf_dim 64, synthetic uniform ids / labels, batch 2048, lr 5e-4, and a
2-class softmax over ``[1, logit]`` as the loss. That package is not in the reference tree, so
NeuMF is restated here from its published definition:

* one embedding table per side holding the MF half (64) and the MLP half (128) of a row;
* GMF branch = elementwise product of the MF halves;
* MLP branch = concat of the MLP halves -> Dense 256 -> 128 -> 64, ReLU each
  (``model_layers[0]`` is the concat width, so the first Dense is 256 -> 256);
* logit = Dense(1) on concat(GMF, MLP).

The reference is fp32-only; here embeddings and activations are fp16 with fp32 master tables
(the CUDA executor's usual mixed precision) and the table gradients are dense
(``--sparse_to_dense_grads`` in the reference's suggested command line).
"""
import torch

from .. import model
from ...nn.builder import NetBuilder
from ...nn.graph import Graph

_NUM_USERS_20M = 138493
_NUM_ITEMS_20M = 26744


class _NcfSyntheticSource:
    """users / items / labels drawn once and reused (reference: local variables,
    T/models/experimental/official_ncf_model.py:99-121)."""

    def __init__(self, batch, num_users, num_items, device, seed, host=False):
        gen = torch.Generator().manual_seed(int(seed))
        self.users = torch.randint(0, num_users, (batch,), generator=gen, dtype=torch.int32)
        self.items = torch.randint(0, num_items, (batch,), generator=gen, dtype=torch.int32)
        self.labels = torch.randint(0, 2, (batch,), generator=gen, dtype=torch.int32)
        dev = torch.device(device)
        if dev.type == "cuda" and not host:
            self.users, self.items, self.labels = (t.to(dev) for t in
                                                   (self.users, self.items, self.labels))
        elif dev.type == "cuda":
            self.users, self.items, self.labels = (t.pin_memory() for t in
                                                   (self.users, self.items, self.labels))
            self.host_resident = True
        self.bytes_per_step = 3 * 4 * batch

    def next_batch(self):
        return self.users, self.items, self.labels


class NcfModel(model.Model):
    def __init__(self, params=None, num_users=_NUM_USERS_20M, num_items=_NUM_ITEMS_20M,
                 model_layers=(256, 256, 128, 64), mf_dim=64):
        super().__init__("official_ncf", batch_size=2048, learning_rate=0.0005,
                         fp16_loss_scale=128, params=params)
        self.num_users, self.num_items = num_users, num_items
        self.model_layers, self.mf_dim = tuple(model_layers), mf_dim
        # no image: these keep the generic benchmark plumbing (header, preprocessing) happy
        self.image_size, self.depth = 1, 1
        self.aux_loss_weight = 0.0

    def get_input_shapes(self, subset="train"):
        del subset
        return [[self.batch_size], [self.batch_size], [self.batch_size]]

    def skip_final_affine_layer(self):
        return True

    def build_graph(self, nclass=None, phase_train=True, batch_size=None) -> Graph:
        n = batch_size or self.batch_size
        g = Graph(n, input_shape=(n,), input_dtype="i32")
        items = g.add_input((n,), "i32")
        cnn = NetBuilder(g, phase_train=phase_train)
        mlp_half = self.model_layers[0] // 2
        dim = self.mf_dim + mlp_half
        mf_user = cnn.embedding(g.input, self.num_users, dim, 0, self.mf_dim,
                                scope_name="embedding_user")
        user_table = mf_user.producer.params["table"].name
        mlp_user = cnn.embedding(g.input, self.num_users, dim, self.mf_dim, mlp_half,
                                 scope_name="embedding_user", share_with=user_table)
        mf_item = cnn.embedding(items, self.num_items, dim, 0, self.mf_dim,
                                scope_name="embedding_item")
        item_table = mf_item.producer.params["table"].name
        mlp_item = cnn.embedding(items, self.num_items, dim, self.mf_dim, mlp_half,
                                 scope_name="embedding_item", share_with=item_table)
        mf = cnn.mul(mf_user, mf_item)
        cnn.concat([mlp_user, mlp_item])
        for width in self.model_layers[1:]:
            cnn.affine(width, activation="relu")
        cnn.concat([mf, cnn.top_layer])
        cnn.affine(1, activation="linear", scope_name="rating")
        g.logits = cnn.top_layer
        g.loss = {"kind": "ncf"}
        return g

    def make_input_source(self, batch, device, seed, rank=0, host=False):
        return _NcfSyntheticSource(batch, self.num_users, self.num_items, device,
                                   seed * 1000003 + rank, host=host)
