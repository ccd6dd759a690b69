from garage_b200.np.policies.scripted_policy import ScriptedPolicy

__all__ = ['ScriptedPolicy']
