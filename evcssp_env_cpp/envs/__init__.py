from charginghub_env_b200.env import BatchedHubEnv, EvcsspManagerEnv_v6  # noqa: F401
