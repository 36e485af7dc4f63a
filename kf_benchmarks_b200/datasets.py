"""Dataset descriptors (T/datasets.py:41-251): name, class count, epoch sizes and whether
inputs are synthetic (no ``data_dir``). Readers for real data live in ``data/``."""
from __future__ import annotations

import os

IMAGENET_NUM_TRAIN_IMAGES = 1281167
IMAGENET_NUM_VAL_IMAGES = 50000
COCO_NUM_TRAIN_IMAGES = 118287
COCO_NUM_VAL_IMAGES = 4952


class Dataset:
    def __init__(self, name, data_dir=None, queue_runner_required=False, num_classes=None):
        self.name = name
        self.data_dir = data_dir
        self._queue_runner_required = queue_runner_required
        self._num_classes = num_classes

    def tf_record_pattern(self, subset):
        return os.path.join(self.data_dir, "%s-*-of-*" % subset)

    @property
    def num_classes(self):
        return self._num_classes

    @num_classes.setter
    def num_classes(self, val):
        self._num_classes = val

    def num_examples_per_epoch(self, subset="train"):
        raise NotImplementedError

    def __str__(self):
        return self.name

    def queue_runner_required(self):
        return self._queue_runner_required

    def use_synthetic_gpu_inputs(self):
        return not self.data_dir

    def get_input_preprocessor(self, input_preprocessor="default"):
        from .data import preprocessing
        return preprocessing.get_preprocessor(self.name, input_preprocessor)


class ImageDataset(Dataset):
    def __init__(self, name, height, width, depth=None, data_dir=None, queue_runner_required=False,
                 num_classes=1001):
        super().__init__(name, data_dir, queue_runner_required, num_classes)
        self.height = height
        self.width = width
        self.depth = depth or 3


class ImagenetDataset(ImageDataset):
    """1001 classes (label 0 is background), T/datasets.py:124-137."""

    def __init__(self, data_dir=None):
        super().__init__("imagenet", 300, 300, data_dir=data_dir)

    def num_examples_per_epoch(self, subset="train"):
        if subset == "train":
            return IMAGENET_NUM_TRAIN_IMAGES
        if subset == "validation":
            return IMAGENET_NUM_VAL_IMAGES
        raise ValueError('Invalid data subset "%s"' % subset)


class Cifar10Dataset(ImageDataset):
    """11 classes in the reference (T/datasets.py:140-153)."""

    def __init__(self, data_dir=None):
        super().__init__("cifar10", 32, 32, data_dir=data_dir, queue_runner_required=True,
                         num_classes=11)

    def read_data_files(self, subset="train"):
        from .data import cifar
        return cifar.read_data_files(self.data_dir, subset)

    def num_examples_per_epoch(self, subset="train"):
        if subset == "train":
            return 50000
        if subset == "validation":
            return 10000
        raise ValueError('Invalid data subset "%s"' % subset)


class COCODataset(ImageDataset):
    def __init__(self, data_dir=None, image_size=300):
        super().__init__("coco", image_size, image_size, data_dir=data_dir, num_classes=81)

    def num_examples_per_epoch(self, subset="train"):
        return COCO_NUM_TRAIN_IMAGES if subset == "train" else COCO_NUM_VAL_IMAGES


class LibrispeechDataset(Dataset):
    def __init__(self, data_dir=None):
        super().__init__("librispeech", data_dir, num_classes=29)

    def num_examples_per_epoch(self, subset="train"):
        return 2


_SUPPORTED_DATASETS = {
    "imagenet": ImagenetDataset,
    "cifar10": Cifar10Dataset,
    "librispeech": LibrispeechDataset,
    "coco": COCODataset,
}


def create_dataset(data_dir, data_name):
    """T/datasets.py:232-251: no data_dir and no name -> synthetic ImageNet."""
    if not data_dir and not data_name:
        data_name = "imagenet"
    if data_name is None:
        for supported in _SUPPORTED_DATASETS:
            if supported in data_dir:
                data_name = supported
                break
        else:
            raise ValueError("Could not identify name of dataset. Please specify with "
                             "--data_name option.")
    if data_name not in _SUPPORTED_DATASETS:
        raise ValueError("Unknown dataset. Must be one of %s" %
                         ", ".join(sorted(_SUPPORTED_DATASETS)))
    return _SUPPORTED_DATASETS[data_name](data_dir)
