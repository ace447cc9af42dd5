from . import ActionWrapper, ActType, Env, ObservationWrapper, ObsType, RewardWrapper, Wrapper  # noqa: F401
