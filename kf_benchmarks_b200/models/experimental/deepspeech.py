"""DeepSpeech2 (``--model=deepspeech2 --data_name=librispeech``).

Reference: T/models/experimental/deepspeech.py:121-442.

* input spectrograms [N, 3494, 161, 1]; two convolutions with explicit padding, no bias,
  ReLU6 then batch norm: 41x11 stride (2,2) pad (20,5) and 21x11 stride (2,1) pad (10,5),
  32 filters each (:296-334) -> [N, 874, 81, 32] -> [N, 874, 2592];
* five bidirectional BasicLSTMCell(800) layers, batch norm in front of layers 2..5
  (:232-268, 341-347); batch norm; dense to 29 classes (:348-350);
* CTC loss on log(softmax + eps) with blank = class 28, input length
  floor(actual * T_ctc / 3494), mean over the batch (:359-396);
* synthetic inputs: uniform spectrogram, labels uniform in [0, 28), full lengths (:283-294);
* evaluation decodes greedily and reports CER / WER (:32-118, 413-442).

Batch norm: momentum .997, eps 1e-5, center+scale (tf.layers defaults), over the last axis.
"""
import numpy as np
import torch

from .. import model
from ...nn.builder import NetBuilder
from ...nn.graph import Graph


class DeepSpeechDecoder:
    """Greedy decoder + CER/WER (T/models/experimental/deepspeech.py:32-118)."""

    def __init__(self, labels, blank_index=28):
        self.labels = labels
        self.blank_index = blank_index
        self.int_to_char = dict(enumerate(labels))

    def convert_to_string(self, sequence):
        return "".join(self.int_to_char[int(i)] for i in sequence)

    @staticmethod
    def _levenshtein(a, b):
        prev = list(range(len(b) + 1))
        for i, ca in enumerate(a, 1):
            cur = [i]
            for j, cb in enumerate(b, 1):
                cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (ca != cb)))
            prev = cur
        return prev[-1]

    def wer(self, decode, target):
        """Word-level edit distance between two strings."""
        return self._levenshtein(decode.split(), target.split())

    def cer(self, decode, target):
        """Character-level edit distance."""
        return self._levenshtein(decode, target)

    def decode(self, char_indexes):
        """Index sequence -> string (used for the ground truth)."""
        return self.convert_to_string(char_indexes)

    def decode_logits(self, logits):
        """Best path: argmax per frame, merge repeats, drop blanks."""
        best = list(np.argmax(np.asarray(logits), axis=1))
        merged = [k for i, k in enumerate(best) if i == 0 or k != best[i - 1]]
        return self.convert_to_string([k for k in merged if k != self.blank_index])


class _SpeechSyntheticSource:
    def __init__(self, batch, t, f, lmax, device, seed, host=False):
        gen = torch.Generator().manual_seed(int(seed))
        self.tensors = [
            torch.rand((batch, t, f, 1), generator=gen).half(),
            torch.randint(0, 28, (batch, lmax), generator=gen, dtype=torch.int32),
            torch.full((batch, 1), t, dtype=torch.int32),
            torch.full((batch, 1), lmax, dtype=torch.int32),
        ]
        dev = torch.device(device)
        if dev.type == "cuda" and not host:
            self.tensors = [x.to(dev) for x in self.tensors]
        elif dev.type == "cuda":
            self.tensors = [x.pin_memory() for x in self.tensors]
            self.host_resident = True
        self.bytes_per_step = sum(x.numel() * x.element_size() for x in self.tensors)

    def next_batch(self):
        return tuple(self.tensors)


class LibrispeechInputSource:
    """Preprocessed LibriSpeech TFRecords (T/preprocessing.py:977-1112): SequenceExamples with
    per-frame 161-bin spectrogram ``features`` and ``labels`` / ``input_length`` /
    ``label_length`` in the context; batches are zero-padded to the model's static shapes
    (``train-clean-*.tfrecords`` / ``test-clean.tfrecords``, T/datasets.py:93-99)."""
    host_resident = True

    def __init__(self, data_dir, batch, t, f, lmax, device, seed, rank=0, world=1, train=True):
        import glob
        import os
        from ...data import tfrecord
        pat = "train-clean-*.tfrecords" if train else "test-clean.tfrecords"
        files = sorted(glob.glob(os.path.join(data_dir, pat)))
        if not files:
            raise FileNotFoundError("Found no files in --data_dir matching: %s" %
                                    os.path.join(data_dir, pat))
        self.files = files[rank::world] or files
        self.batch, self.t, self.f, self.lmax, self.train = batch, t, f, lmax, train
        self.rng = np.random.RandomState(seed * 7919 + rank)
        self._tf = tfrecord
        self._it = self._records()
        self.pin = torch.device(device).type == "cuda"
        self.bytes_per_step = batch * (t * f * 2 + lmax * 4 + 8)

    def _records(self):
        while True:
            order = list(self.files)
            if self.train:
                self.rng.shuffle(order)
            for path in order:
                for rec in self._tf.read_records(path):
                    yield rec
            if not self.train:
                return

    def next_batch(self):
        x = np.zeros((self.batch, self.t, self.f, 1), np.float32)
        lab = np.zeros((self.batch, self.lmax), np.int32)
        ilen = np.zeros((self.batch, 1), np.int32)
        llen = np.zeros((self.batch, 1), np.int32)
        n = 0
        while n < self.batch:
            try:
                rec = next(self._it)
            except StopIteration:
                if n == 0:
                    raise
                break
            ctx, lists = self._tf.parse_sequence_example(rec)
            feats = np.asarray(lists.get("features", []), np.float32).reshape(-1, self.f)
            steps = min(len(feats), self.t)
            x[n, :steps, :, 0] = feats[:steps]
            labels = np.asarray(ctx.get("labels", []), np.int32)[:self.lmax]
            lab[n, :len(labels)] = labels
            ilen[n, 0] = min(int(ctx.get("input_length", [steps])[0]), self.t)
            llen[n, 0] = min(int(ctx.get("label_length", [len(labels)])[0]), self.lmax)
            n += 1
        out = [torch.from_numpy(x).half(), torch.from_numpy(lab), torch.from_numpy(ilen),
               torch.from_numpy(llen)]
        if self.pin:
            out = [v.pin_memory() for v in out]
        return tuple(out)


class DeepSpeech2Model(model.Model):
    BATCH_NORM_EPSILON = 1e-5
    BATCH_NORM_DECAY = 0.997
    CONV_FILTERS = 32

    def __init__(self, num_rnn_layers=5, rnn_type="lstm", is_bidirectional=True,
                 rnn_hidden_size=800, use_bias=True, params=None):
        super().__init__("deepspeech2", batch_size=128, learning_rate=0.0005,
                         fp16_loss_scale=128, params=params)
        if rnn_type != "lstm":
            raise NotImplementedError("only the LSTM cell (the reference default) is built")
        if not use_bias:
            raise NotImplementedError("the final dense layer is built with its bias")
        self.num_rnn_layers = num_rnn_layers
        self.rnn_type = rnn_type
        self.is_bidirectional = is_bidirectional
        self.rnn_hidden_size = rnn_hidden_size
        self.use_bias = use_bias
        self.num_feature_bins = 161
        self.max_time_steps = 3494
        self.max_label_length = 576
        self.image_size, self.depth = 1, 1
        self.aux_loss_weight = 0.0

    def get_input_shapes(self, subset="train"):
        del subset
        return [[self.batch_size, self.max_time_steps, self.num_feature_bins, 1],
                [self.batch_size, self.max_label_length], [self.batch_size, 1],
                [self.batch_size, 1]]

    def skip_final_affine_layer(self):
        return True

    def _bn(self, cnn):
        return cnn.batch_norm(decay=self.BATCH_NORM_DECAY, epsilon=self.BATCH_NORM_EPSILON,
                              scale=True)

    def build_graph(self, nclass=29, phase_train=True, batch_size=None) -> Graph:
        n = batch_size or self.batch_size
        nclass = nclass or 29
        g = Graph(n, input_shape=(n, self.max_time_steps, self.num_feature_bins, 1))
        cnn = NetBuilder(g, phase_train=phase_train)
        cnn.conv(self.CONV_FILTERS, 41, 11, 2, 2, explicit_pad=(20, 5), use_batch_norm=False,
                 activation="relu6", bias=None, scope_name="cnn_1")
        self._bn(cnn)
        cnn.conv(self.CONV_FILTERS, 21, 11, 2, 1, explicit_pad=(10, 5), use_batch_norm=False,
                 activation="relu6", bias=None, scope_name="cnn_2")
        self._bn(cnn)
        cnn.merge_last_dims()
        for layer in range(self.num_rnn_layers):
            if layer != 0:
                self._bn(cnn)
            cnn.lstm(self.rnn_hidden_size, self.is_bidirectional,
                     scope_name="rnn_%d" % (layer + 1))
        self._bn(cnn)
        cnn.affine(nclass, activation="linear", scope_name="dense")
        g.logits = cnn.top_layer
        g.label_specs = [((n, self.max_label_length), "i32"), ((n, 1), "i32"), ((n, 1), "i32")]
        g.loss = {"kind": "ctc", "max_time_steps": self.max_time_steps, "eps": 1e-7}
        return g

    def make_input_source(self, batch, device, seed, rank=0, host=False, world=1, train=True):
        if self.params is not None and self.params.data_dir:
            return LibrispeechInputSource(self.params.data_dir, batch, self.max_time_steps,
                                          self.num_feature_bins, self.max_label_length, device,
                                          seed, rank, world, train)
        return _SpeechSyntheticSource(batch, self.max_time_steps, self.num_feature_bins,
                                      self.max_label_length, device, seed * 1000003 + rank,
                                      host=host)

    # -- evaluation: CTC loss + greedy decoding, CER / WER at the end ------------------------
    def eval_step(self, executor, batch, src):
        executor.set_inputs(*batch)
        executor.forward(False)
        executor.loss_backward(backward=False)
        loss = executor.read_loss()[0]
        probs = torch.softmax(executor.logits().float(), dim=-1).cpu().numpy()
        labels, _, lab_len = (t.cpu().numpy() for t in batch[1:])
        if not hasattr(self, "_eval_items"):
            self._eval_items = {"probs": [], "labels": []}
        for i in range(probs.shape[0]):
            self._eval_items["probs"].append(probs[i])
            self._eval_items["labels"].append(labels[i][:int(lab_len[i, 0])])
        return loss

    def eval_finish(self, params):
        items = getattr(self, "_eval_items", None)
        self._eval_items = {"probs": [], "labels": []}
        if not items or not items["probs"]:
            return {}
        return self.postprocess(items)

    def postprocess(self, results):
        """results: {'probs': [N, T, 29], 'labels': [N, L]} -> logs CER / WER
        (T/models/experimental/deepspeech.py:413-442)."""
        from ...utils import log_util
        probs, targets = results["probs"], results["labels"]
        speech_labels = " abcdefghijklmnopqrstuvwxyz'-"
        dec = DeepSpeechDecoder(speech_labels)
        total_cer = total_wer = 0.0
        n = len(probs)
        for i in range(n):
            predicted = dec.decode_logits(probs[i])
            expected = dec.decode(targets[i])
            total_cer += dec.cer(predicted, expected) / float(max(len(expected), 1))
            total_wer += dec.wer(predicted, expected) / float(max(len(expected.split()), 1))
        total_cer /= n
        total_wer /= n
        log_util.log_fn("total CER: {:f}; total WER: {:f}; total example: {:d}.".format(
            total_cer, total_wer, n))
        return {"top_1_accuracy": 0.0, "top_5_accuracy": 0.0, "cer": total_cer, "wer": total_wer}
