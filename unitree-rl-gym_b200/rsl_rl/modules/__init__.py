from .actor_critic import ActorCritic  # noqa: F401
from .actor_critic_recurrent import ActorCriticRecurrent, Memory  # noqa: F401
