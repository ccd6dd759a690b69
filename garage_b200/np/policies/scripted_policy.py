"""ScriptedPolicy (reference np/policies/scripted_policy.py:6-60): the action is looked up by the
observation.  The table is also what `b2g_gridworld_rollout` reads on the device."""
import numpy as np


class ScriptedPolicy:

    def __init__(self, scripted_actions, agent_env_infos=None):
        self._scripted_actions = scripted_actions
        self._agent_env_infos = agent_env_infos
        self.name = 'ScriptedPolicy'

    def set_param_values(self, params):
        """No parameters."""

    def get_param_values(self):
        return []

    def reset(self, do_resets=None):
        """Stateless."""

    def action_table(self, n_states):
        """int32 array of length n_states (list or dict tables alike)."""
        return np.asarray([int(self._scripted_actions[s]) for s in range(n_states)], dtype=np.int32)

    def get_action(self, observation):
        info = self._agent_env_infos[observation] if self._agent_env_infos else dict()
        return self._scripted_actions[observation], info

    def get_actions(self, observations):
        info = self._agent_env_infos[observations[0]] if self._agent_env_infos else dict()
        return [self._scripted_actions[obs] for obs in observations], info
