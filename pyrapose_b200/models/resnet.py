"""ResNet backbone entry points — mirror of PyraPose/models/resnet.py (ResNetBackbone :29-76,
resnet_retinanet :79-110). keras-resnet's ResNet50 graph is executed by the native engine; the ImageNet
checkpoint download (:89-98) is replaced by random initialisation (no network) — load real weights with
model.load_weights()."""
from . import Backbone
from . import retinanet
from ..utils.image import preprocess_image


class ResNetBackbone(Backbone):
    """ Describes backbone information and provides utility functions."""

    def retinanet(self, *args, **kwargs):
        """ Returns a retinanet model using the correct backbone."""
        return resnet_retinanet(*args, backbone=self.backbone, **kwargs)

    def download_imagenet(self):
        raise RuntimeError("no network access: provide ResNet-50 weights with model.load_weights(path)")

    def validate(self):
        """ Checks whether the backbone string is correct."""
        allowed_backbones = ['resnet50', 'resnet101', 'resnet152']
        backbone = self.backbone.split('_')[0]
        if backbone not in allowed_backbones:
            raise ValueError('Backbone (\'{}\') not in allowed backbones ({}).'.format(backbone, allowed_backbones))

    def preprocess_image(self, inputs):
        """ Takes as input an image and prepares it for being passed through the network."""
        return preprocess_image(inputs, mode='caffe')


def resnet_retinanet(num_classes, backbone='resnet50', inputs=None, modifier=None, **kwargs):
    """ Constructs a PFPN model on a ResNet-50 backbone. The reference always builds ResNet50 here regardless of
    the backbone string (resnet.py:87); deeper variants are rejected explicitly instead."""
    if backbone.split('_')[0] != 'resnet50':
        raise NotImplementedError("only the ResNet-50 backbone is on the hot path (resnet.py:87 builds ResNet50)")
    if modifier is not None:
        raise NotImplementedError("backbone modifiers operate on Keras graphs; not supported")
    return retinanet.retinanet(inputs=inputs, num_classes=num_classes, backbone_layers=None, **kwargs)


def resnet50_retinanet(num_classes, inputs=None, **kwargs):
    return resnet_retinanet(num_classes=num_classes, backbone='resnet50', inputs=inputs, **kwargs)
