"""Drop-in module name of the reference package (evcssp_env_cpp/__init__.py:1-8).

`gym.make('evcssp_env_cpp:charging-hub-v6', **env_kwargs)` (test/env_test.py:34) resolves here when gym or
gymnasium is installed; without either, import the class directly:
    from evcssp_env_cpp.envs import EvcsspManagerEnv_v6
The environment is stepped by the B200 CUDA engine (charginghub_env_b200); there is no CPU path.
"""
from charginghub_env_b200.env import register as _register

_register()
