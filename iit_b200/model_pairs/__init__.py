from iit_b200.model_pairs.nodes import HLNode, LLNode
from iit_b200.model_pairs.base_model_pair import BaseModelPair
from iit_b200.model_pairs.iit_model_pair import IITModelPair
from iit_b200.model_pairs.iit_behavior_model_pair import IITBehaviorModelPair
from iit_b200.model_pairs.strict_iit_model_pair import StrictIITModelPair
from iit_b200.model_pairs.ioi_model_pair import IOI_ModelPair
from iit_b200.model_pairs.freeze_model_pair import FreezedModelPair
from iit_b200.model_pairs.stop_grad_pair import StopGradHookedModel, StopGradModelPair

__all__ = ["HLNode", "LLNode", "BaseModelPair", "IITModelPair", "IITBehaviorModelPair", "StrictIITModelPair",
           "IOI_ModelPair", "FreezedModelPair", "StopGradHookedModel", "StopGradModelPair"]
