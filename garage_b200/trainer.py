"""Epoch loop of the on-policy path (mirror of the reference's `Trainer`,
trainer.py:66-537: `setup`, `train`, `step_epochs`, `obtain_episodes`, `step_itr`,
`total_env_steps`).  Snapshotting / dowel output are host services outside the
accelerated path; a `snapshot_dir` enables a cloudpickle snapshot per epoch
(trainer.py:263-293) and `resume`-style restoring via `Trainer.restore`."""
import os
import time

from garage_b200.experiment import deterministic
from garage_b200.experiment.snapshotter import Snapshotter, load_snapshot
from garage_b200.tabular import tabular


class TrainArgs:
    """Reference trainer.py:33-63."""

    def __init__(self, n_epochs, batch_size, plot, store_episodes, pause_for_plot, start_epoch):
        self.n_epochs = n_epochs
        self.batch_size = batch_size
        self.plot = plot
        self.store_episodes = store_episodes
        self.pause_for_plot = pause_for_plot
        self.start_epoch = start_epoch


class NotSetupError(Exception):
    """Reference trainer.py:540."""


class Trainer:
    """Sets up and runs the algorithm's epoch loop.

    Args:
        snapshot_config: object with `snapshot_dir`, `snapshot_mode`, `snapshot_gap` (the
            reference's `SnapshotConfig`), or None for no snapshots.
    """

    def __init__(self, snapshot_config=None):
        self._snapshot_config = snapshot_config
        # reference trainer.py:106-109; no config = no snapshots (the reference always has one)
        self._snapshotter = None if snapshot_config is None else Snapshotter(
            snapshot_config.snapshot_dir, snapshot_config.snapshot_mode,
            snapshot_config.snapshot_gap)
        self._has_setup = False
        self._seed = None
        self._train_args = None
        self._stats_total_env_steps = 0
        self._stats_total_epoch = 0
        self._stats_total_itr = 0
        self._stats_last_episode = None
        self._start_time = None
        self._itr_start_time = None
        self.step_itr = None
        self.step_episode = None
        self.enable_logging = True
        self._algo = None
        self._env = None
        self._sampler = None
        self.log = []  # one dict of tabular keys per epoch (dowel is not available here)

    @property
    def total_env_steps(self):
        return self._stats_total_env_steps

    @total_env_steps.setter
    def total_env_steps(self, value):
        self._stats_total_env_steps = value

    def setup(self, algo, env):
        """Reference trainer.py:137-162: the sampler is `algo._sampler`."""
        self._algo = algo
        self._env = env
        self._seed = deterministic.get_seed()
        if hasattr(algo, '_sampler'):
            self._sampler = algo._sampler
        self._has_setup = True

    def _shutdown_worker(self):
        if self._sampler is not None:
            self._sampler.shutdown_worker()

    def obtain_episodes(self, itr, batch_size=None, agent_update=None, env_update=None):
        """Reference trainer.py:179-229: agent_update defaults to the policy's parameters."""
        if self._sampler is None:
            raise ValueError('trainer was not initialized with `sampler`. the algo should have a '
                             '`_sampler` field when `setup()` is called')
        if batch_size is None and self._train_args.batch_size is None:
            raise ValueError('trainer was not initialized with `batch_size`. Either provide '
                             '`batch_size` to trainer.train,  or pass `batch_size` to '
                             'trainer.obtain_samples.')
        if agent_update is None:
            policy = getattr(self._algo, 'exploration_policy', None) or self._algo.policy
            # the sampler shares the policy object: its flat CUDA parameter vector IS the update
            agent_update = policy if getattr(self._sampler, '_policy', None) is policy \
                else policy.get_param_values()
        episodes = self._sampler.obtain_samples(
            itr, (batch_size or self._train_args.batch_size), agent_update=agent_update,
            env_update=env_update)
        self._stats_total_env_steps += getattr(episodes, 'n_steps', None) or int(sum(episodes.lengths))
        return episodes

    def obtain_samples(self, itr, batch_size=None, agent_update=None, env_update=None):
        eps = self.obtain_episodes(itr, batch_size, agent_update, env_update)
        self._stats_last_episode = eps
        return eps.to_list()

    def save(self, epoch):
        """Reference trainer.py:263-293: seed, train_args, stats, env and algo through the
        Snapshotter (cloudpickle; device buffers are dropped by the classes' __getstate__)."""
        if not self._has_setup:
            raise NotSetupError('Use setup() to setup trainer before saving.')
        if self._snapshotter is None:
            return
        params = dict(seed=self._seed, train_args=self._train_args,
                      stats=dict(total_epoch=self._stats_total_epoch, total_itr=self._stats_total_itr,
                                 total_env_steps=self._stats_total_env_steps,
                                 last_episode=self._stats_last_episode),
                      env=self._env, algo=self._algo)
        self._snapshotter.save_itr_params(epoch, params)

    def restore(self, from_dir, from_epoch='last'):
        """Reference trainer.py:295-341: load a snapshot, re-seed, set the trainer up with the
        saved env and algo, continue at the epoch after the saved one.  Returns the TrainArgs."""
        saved = load_snapshot(from_dir, from_epoch)
        self._seed = saved['seed']
        self._train_args = saved['train_args']
        st = saved['stats']
        self._stats_total_epoch, self._stats_total_itr = st['total_epoch'], st['total_itr']
        self._stats_total_env_steps = st['total_env_steps']
        self._stats_last_episode = st.get('last_episode')
        if self._seed is not None:
            deterministic.set_seed(self._seed)
        self.setup(env=saved['env'], algo=saved['algo'])
        self._train_args.start_epoch = self._stats_total_epoch + 1
        return self._train_args

    def log_diagnostics(self, pause_for_plot=False):
        tabular.record('TotalEnvSteps', self._stats_total_env_steps)
        tabular.record('Time', time.time() - self._start_time)
        tabular.record('EpochTime', time.time() - self._itr_start_time)
        self.log.append(tabular.as_dict)
        tabular.dump()

    def train(self, n_epochs, batch_size=None, plot=False, store_episodes=False,
              pause_for_plot=False):
        """Reference trainer.py:361-405."""
        if not self._has_setup:
            raise NotSetupError('Use setup() to setup trainer before training.')
        if plot:
            raise NotImplementedError('plotting is unavailable on this box (no graphics stack)')
        self._train_args = TrainArgs(n_epochs=n_epochs, batch_size=batch_size, plot=plot,
                                     store_episodes=store_episodes, pause_for_plot=pause_for_plot,
                                     start_epoch=0)
        average_return = self._algo.train(self)
        self._shutdown_worker()
        return average_return

    def step_epochs(self):
        """Generator over epochs (reference trainer.py:407-455): after each epoch body
        save a snapshot and dump the diagnostics."""
        self._start_time = time.time()
        self.step_itr = self._stats_total_itr
        self.step_episode = None
        n_epochs = self._train_args.n_epochs
        if 'GARAGE_EXAMPLE_TEST_N_EPOCHS' in os.environ:  # trainer.py:433-435
            n_epochs = int(os.environ['GARAGE_EXAMPLE_TEST_N_EPOCHS'])
        for epoch in range(self._train_args.start_epoch, n_epochs):
            self._itr_start_time = time.time()
            tabular.clear()
            yield epoch
            self._stats_last_episode = self.step_episode if self._train_args.store_episodes else None  # trainer.py:443-446
            self._stats_total_epoch = epoch
            self._stats_total_itr = self.step_itr
            self.save(epoch)
            if self.enable_logging:
                self.log_diagnostics(self._train_args.pause_for_plot)

    def resume(self, n_epochs=None, batch_size=None):
        """Reference trainer.py:457-500."""
        if self._train_args is None:
            raise NotSetupError('You must restore a checkpoint before resume.')
        self._train_args.n_epochs = n_epochs or self._train_args.n_epochs
        self._train_args.batch_size = batch_size or self._train_args.batch_size
        average_return = self._algo.train(self)
        self._shutdown_worker()
        return average_return
