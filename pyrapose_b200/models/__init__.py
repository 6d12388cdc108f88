"""Model API — mirror of PyraPose/models/__init__.py (Backbone :6-52, backbone() :55-65, load_model :68-71,
convert_model :74-76, assert_training_model :79-81, check_training_model :84-89)."""
from __future__ import print_function


class Backbone(object):
    """ This class stores additional information on backbones."""

    def __init__(self, backbone):
        from .. import layers
        from .. import initializers
        # names a serialized reference model refers to -> the objects that implement them here
        self.custom_objects = {
            'UpsampleLike': layers.UpsampleLike,
            'PriorProbability': initializers.PriorProbability,
            'RegressBoxes': layers.RegressBoxes,
            'FilterDetections': layers.FilterDetections,
            'Anchors': layers.Anchors,
            'ClipBoxes': layers.ClipBoxes,
            'RegressBoxes3D': layers.RegressBoxes3D,
            'DenormBoxes3D': layers.DenormBoxes3D,
        }
        self.backbone = backbone
        self.validate()

    def retinanet(self, *args, **kwargs):
        raise NotImplementedError('retinanet method not implemented.')

    def validate(self):
        raise NotImplementedError('validate method not implemented.')

    def preprocess_image(self, inputs):
        raise NotImplementedError('preprocess_image method not implemented.')


def backbone(backbone_name):
    """ Returns a backbone object for the given backbone name."""
    if 'resnet' in backbone_name:
        from .resnet import ResNetBackbone as b
    else:
        # the reference's densenet / efficientnet backbones are stale (their retinanet() call no longer matches
        # retinanet.retinanet's signature, PyraPose/models/densenet.py:100-101) and are out of the hot path
        raise NotImplementedError('Backbone class for  \'{}\' not implemented.'.format(backbone_name))
    return b(backbone_name)


def load_model(filepath, backbone_name='resnet50'):
    """ Loads a model saved with `model.save(filepath)`: a Keras HDF5 snapshot (this package's or the reference's
    ModelCheckpoint files, read by utils/hdf5.py and mapped by models/keras_io.py) or an npz container keyed by Keras
    layer names (HWIO kernels)."""
    from .retinanet import PFPNModel
    backbone(backbone_name)  # validates the name like the reference does
    return PFPNModel.load(filepath)


def convert_model(model, nms=True, class_specific_filter=True, anchor_params=None, precision=None, alias_arena=False):
    """ Converts a training model to an inference model (adds anchors + RegressBoxes3D). precision=None keeps the
    training model's ('bf16'); 'fp32' builds the 1e-4 path (six MMAs over three-part bf16 splits) on the same weights."""
    from .retinanet import retinanet_bbox
    return retinanet_bbox(model=model, nms=nms, class_specific_filter=class_specific_filter,
                          anchor_params=anchor_params, precision=precision, alias_arena=alias_arena)


def assert_training_model(model):
    assert (all(output in model.output_names for output in ['3Dbox', 'cls', 'mask'])), \
        "Input is not a training model. Outputs were found, outputs are: {}).".format(model.output_names)


def check_training_model(model):
    """ Like the reference, but raises instead of sys.exit(1) (a library must not exit the process)."""
    try:
        assert_training_model(model)
    except AssertionError as e:
        raise ValueError(str(e))
