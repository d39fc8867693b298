from .impala import ImpalaLearner  # noqa: F401
from .ppo import PPOLearner  # noqa: F401
