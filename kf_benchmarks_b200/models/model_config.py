"""Model registry (T/models/model_config.py:38-142). Unlike the live fork, which only
registers the ResNets, every CNN of the zoo is registered here."""
from __future__ import annotations


from .experimental import deepspeech, official_ncf_model
from . import ssd_model
from . import (alexnet_model, densenet_model, googlenet_model, inception_model, lenet_model,
               mobilenet_v2, nasnet_model,
               official_resnet_model, overfeat_model, resnet_model, trivial_model, vgg_model)

_model_name_to_imagenet_model = {
    "vgg11": vgg_model.Vgg11Model,
    "vgg16": vgg_model.Vgg16Model,
    "vgg19": vgg_model.Vgg19Model,
    "lenet": lenet_model.Lenet5Model,
    "googlenet": googlenet_model.GooglenetModel,
    "overfeat": overfeat_model.OverfeatModel,
    "alexnet": alexnet_model.AlexnetModel,
    "trivial": trivial_model.TrivialModel,
    "inception3": inception_model.Inceptionv3Model,
    "inception4": inception_model.Inceptionv4Model,
    "mobilenet": mobilenet_v2.MobilenetModel,
    "nasnet": nasnet_model.NasnetModel,
    "nasnetlarge": nasnet_model.NasnetLargeModel,
    "ncf": official_ncf_model.NcfModel,
}
_model_name_to_imagenet_model.update(resnet_model.imagenet_factories())
_model_name_to_imagenet_model.update(official_resnet_model.factories())

_model_name_to_cifar_model = {
    "alexnet": alexnet_model.AlexnetCifar10Model,
    "trivial": trivial_model.TrivialCifar10Model,
    "nasnet": nasnet_model.NasnetCifarModel,
    "densenet40_k12": densenet_model.create_densenet40_k12_model,
    "densenet100_k12": densenet_model.create_densenet100_k12_model,
    "densenet100_k24": densenet_model.create_densenet100_k24_model,
}
_model_name_to_cifar_model.update(resnet_model.cifar_factories())

# Every model of the reference's zoo is registered; this map is kept for
# names that may have to be rejected with an explanation in a future build.
_UNSUPPORTED = {}


_model_name_to_object_detection_model = {"ssd300": ssd_model.SSD300Model}
_model_name_to_speech_model = {"deepspeech2": deepspeech.DeepSpeech2Model}


def _get_model_map(dataset_name):
    if dataset_name == "cifar10":
        return _model_name_to_cifar_model
    if dataset_name == "coco":
        return _model_name_to_object_detection_model
    if dataset_name == "librispeech":
        return _model_name_to_speech_model
    if dataset_name in ("imagenet", "synthetic"):
        return _model_name_to_imagenet_model
    raise ValueError("Invalid dataset name: %s" % dataset_name)


def get_model_config(model_name, dataset, params):
    """T/models/model_config.py:126-133."""
    model_map = _get_model_map(dataset.name)
    if model_name not in model_map:
        if model_name in _UNSUPPORTED:
            raise ValueError("Model '%s' is not supported by this build: %s" %
                             (model_name, _UNSUPPORTED[model_name]))
        raise ValueError("Invalid model name '%s' for dataset '%s'" % (model_name, dataset.name))
    return model_map[model_name](params=params)


def register_model(model_name, dataset_name, model_func):
    """T/models/model_config.py:136-142."""
    model_map = _get_model_map(dataset_name)
    if model_name in model_map:
        raise ValueError('Model "%s" is already registered for dataset "%s"' %
                         (model_name, dataset_name))
    model_map[model_name] = model_func


def list_models(dataset_name="imagenet"):
    return sorted(_get_model_map(dataset_name))
