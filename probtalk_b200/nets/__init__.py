"""Mirror of the reference's ``nets`` package for the hot path.

``getattr(probtalk_b200.nets, model_name)(args, config)`` resolves the same three names the reference
registers for this path in nets/__init__.py:5-7."""
from .wrappers import (s2g_body_vqt, s2g_body_predictor, s2g_body_refiner, QuantizationOutput, VQVAE,   # noqa: F401
                       VectorQuantizerEMA)
from .ai1 import All_In_One_Model, init_model   # noqa: F401

_DROP_IN = ("s2g_body_predictor", "s2g_body_refiner", "s2g_body_vqt")


def install_into(reference_nets):
    """Register the drop-in wrappers on the reference's ``nets`` package.

    The reference resolves models with ``getattr(nets, model_name)(args, config)`` (nets/init_model.py:8,
    nets/ai1.py:11), so replacing the three registry names is all that ``scripts/test_holisticbody.py`` /
    ``scripts/demo.py`` need -- no edit of the reference tree:

        import nets, probtalk_b200.nets
        previous = probtalk_b200.nets.install_into(nets)

    Returns the classes that were registered before (``uninstall(nets, previous)`` puts them back)."""
    import sys
    previous = {}
    me = sys.modules[__name__]
    for name in _DROP_IN:
        previous[name] = getattr(reference_nets, name, None)
        setattr(reference_nets, name, getattr(me, name))
    return previous


def uninstall(reference_nets, previous):
    for name, cls in previous.items():
        if cls is None:
            if hasattr(reference_nets, name):
                delattr(reference_nets, name)
        else:
            setattr(reference_nets, name, cls)
