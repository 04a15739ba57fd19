"""Mirror of the reference's ``nets`` package for the hot path.

``getattr(probtalk_b200.nets, model_name)(args, config)`` resolves the same three names the reference
registers for this path in nets/__init__.py:5-7."""
from .wrappers import s2g_body_vqt, s2g_body_predictor, s2g_body_refiner, QuantizationOutput, VQVAE   # noqa: F401
from .ai1 import All_In_One_Model, init_model   # noqa: F401
