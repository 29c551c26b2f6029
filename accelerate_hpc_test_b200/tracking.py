"""Experiment trackers behind ``Accelerator.init_trackers / log / get_tracker / end_training``.

The contract every tracker obeys (identical to the reference's ``GeneralTracker``, tracking.py:101-179):

* class attributes ``name`` (what ``log_with=`` and ``get_tracker`` use), ``requires_logging_directory`` and
  ``main_process_only`` (default ``True``: every method decorated with ``on_main_process`` is a no-op elsewhere);
* ``start()``, ``store_init_configuration(dict)``, ``log(dict, step, **kw)``, ``finish()`` and the ``tracker`` property that
  exposes the third-party run object.

The nine third-party integrations (TensorBoard, Weights & Biases, Trackio, Comet, Aim, MLflow, ClearML, DVCLive, SwanLab)
are thin subclasses of one ``_BackendTracker``: what differs between them — how a run is opened, how the configuration,
scalars and media are handed over, how the run is closed — is written as small per-backend functions, so the control flow
exists once.  ``JSONLTracker`` (new in this build) needs no package: one JSON object per ``log`` call, also the sink of the
built-in performance counters (tokens/s, step ms, exposed-communication ms).
"""

from __future__ import annotations

import json
import os
import time
from functools import wraps
from typing import Any, Optional, Union

from .logging import get_logger
from .state import PartialState
from .utils import imports as _imports
from .utils.dataclasses import LoggerType

logger = get_logger(__name__)


# =========================================================================== plumbing
def on_main_process(function):
    """Decorator for tracker methods: skip the call on non-main processes when the tracker is ``main_process_only``."""

    @wraps(function)
    def maybe_run(self, *args, **kwargs):
        if getattr(self, "main_process_only", False):
            return PartialState().on_main_process(function)(self, *args, **kwargs)
        return function(self, *args, **kwargs)

    return maybe_run


def _scalarize(values: dict) -> dict:
    """Tensors / numpy scalars → Python numbers, arrays → lists (for JSON and for back-ends that want plain numbers)."""
    plain = {}
    for key, value in values.items():
        if hasattr(value, "item") and getattr(value, "numel", lambda: 2)() == 1:
            value = value.item()
        elif hasattr(value, "tolist"):
            value = value.tolist()
        plain[key] = value
    return plain


class GeneralTracker:
    """Base class.  A subclass must define ``name``, ``requires_logging_directory`` and a ``tracker`` property; it usually
    overrides ``store_init_configuration`` and ``log``."""

    main_process_only = True
    _REQUIRED = ("name", "requires_logging_directory", "tracker")

    def __init__(self, _blank: bool = False):
        if _blank:
            return
        have = set(dir(self))
        missing = [f"`{attr}`" for attr in self._REQUIRED if attr not in have]
        if missing:
            raise NotImplementedError(
                "The implementation for this tracker class is missing the following required attributes. Please define them in "
                f"the class definition: {', '.join(missing)}"
            )

    def start(self):
        """Open the run (called lazily by ``Accelerator.init_trackers``)."""

    def store_init_configuration(self, values: dict):
        """Record hyper-parameters once."""

    def log(self, values: dict, step: Optional[int], **kwargs):
        """Record one set of metrics."""

    def finish(self):
        """Close the run."""


# =========================================================================== the dependency-free tracker
class JSONLTracker(GeneralTracker):
    """``{logging_dir}/{run_name}/metrics.jsonl`` (one object per ``log``) and ``hparams.json``."""

    name = "jsonl"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Union[str, os.PathLike], **kwargs):
        super().__init__()
        self.run_name = run_name
        self.logging_dir = os.path.join(str(logging_dir), run_name)
        self._fh = None

    @property
    def tracker(self):
        return self._fh

    @on_main_process
    def start(self):
        os.makedirs(self.logging_dir, exist_ok=True)
        self._fh = open(os.path.join(self.logging_dir, "metrics.jsonl"), "a", buffering=1)

    @on_main_process
    def store_init_configuration(self, values: dict):
        os.makedirs(self.logging_dir, exist_ok=True)
        with open(os.path.join(self.logging_dir, "hparams.json"), "w") as fh:
            json.dump(_scalarize(values), fh, indent=1, default=str)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        if self._fh is None:
            self.start()
        record = {"step": step, "time": time.time(), **_scalarize(values)}
        self._fh.write(json.dumps(record, default=str) + "\n")

    @on_main_process
    def finish(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None


# =========================================================================== third-party back-ends
class _BackendTracker(GeneralTracker):
    """Shared life-cycle of the package-backed trackers: remember the constructor arguments, open the run in ``start``,
    expose it as ``tracker``.  Subclasses implement ``_open`` / ``_config`` / ``_scalars`` / ``_close``."""

    requires_logging_directory = False

    def __init__(self, run_name: Optional[str] = None, logging_dir: Optional[Union[str, os.PathLike]] = None, **kwargs):
        super().__init__()
        self.run_name = run_name
        self.logging_dir = logging_dir
        self.init_kwargs = kwargs
        self._run = None

    @property
    def tracker(self):
        return self._run

    @on_main_process
    def start(self):
        self._run = self._open()
        logger.debug(f"Initialized {self.name} project {self.run_name}")
        logger.debug("Make sure to log any initial configurations with `self.store_init_configuration` before training!")

    @on_main_process
    def store_init_configuration(self, values: dict):
        self._config(values)
        logger.debug(f"Stored initial configuration hyperparameters to {self.name}")

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        self._scalars(values, step, **kwargs)
        logger.debug(f"Successfully logged to {self.name}")

    @on_main_process
    def finish(self):
        self._close()
        logger.debug(f"{self.name} run closed")

    # defaults that fit the "run object with .config / .log / .finish" family (wandb, trackio, swanlab)
    def _config(self, values):
        self._run.config.update(values)

    def _scalars(self, values, step, **kwargs):
        self._run.log(values, step=step, **kwargs)

    def _close(self):
        self._run.finish()


class TensorBoardTracker(_BackendTracker):
    """``torch.utils.tensorboard`` (or tensorboardX) ``SummaryWriter`` under ``{logging_dir}/{run_name}``."""

    name = "tensorboard"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Union[str, os.PathLike], **kwargs):
        super().__init__(run_name, os.path.join(str(logging_dir), run_name), **kwargs)

    def _open(self):
        try:
            from torch.utils import tensorboard
        except ModuleNotFoundError:
            import tensorboardX as tensorboard
        return tensorboard.SummaryWriter(self.logging_dir, **self.init_kwargs)

    def _config(self, values):
        self._run.add_hparams(values, metric_dict={})
        self._run.flush()
        # a human-readable copy next to the event files, one folder per launch
        stamp_dir = os.path.join(self.logging_dir, str(time.time()))
        os.makedirs(stamp_dir)
        import yaml

        with open(os.path.join(stamp_dir, "hparams.yml"), "w") as fh:
            try:
                yaml.dump(values, fh)
            except yaml.representer.RepresenterError:
                logger.error("Serialization to store hyperparameters failed")
                raise

    def _scalars(self, values, step, **kwargs):
        writer = self._run
        for key, value in _scalarize(values).items():
            if isinstance(value, (int, float)):
                writer.add_scalar(key, value, global_step=step, **kwargs)
            elif isinstance(value, str):
                writer.add_text(key, value, global_step=step, **kwargs)
            elif isinstance(value, dict):
                writer.add_scalars(key, value, global_step=step, **kwargs)
        writer.flush()

    @on_main_process
    def log_images(self, values: dict, step: Optional[int], **kwargs):
        """``values``: tag → ``[N, C, H, W]`` batch of images."""
        for tag, batch in values.items():
            self._run.add_images(tag, batch, global_step=step, **kwargs)

    def _close(self):
        self._run.close()


class WandBTracker(_BackendTracker):
    name = "wandb"
    main_process_only = True

    def _open(self):
        import wandb

        return wandb.init(project=self.run_name, **self.init_kwargs)

    def _config(self, values):
        self._run.config.update(values, allow_val_change=True)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        import wandb

        for tag, images in values.items():
            self._run.log({tag: [wandb.Image(img) for img in images]}, step=step, **kwargs)

    @on_main_process
    def log_table(self, table_name: str, columns=None, data=None, dataframe=None, step: Optional[int] = None, **kwargs):
        import wandb

        self._run.log({table_name: wandb.Table(columns=columns, data=data, dataframe=dataframe)}, step=step, **kwargs)


class TrackioTracker(_BackendTracker):
    name = "trackio"

    def _open(self):
        import trackio

        return trackio.init(project=self.run_name, **self.init_kwargs)

    def _scalars(self, values, step, **kwargs):
        self._run.log(values, **kwargs)


class SwanLabTracker(_BackendTracker):
    name = "swanlab"

    def _open(self):
        import swanlab

        run = swanlab.init(project=self.run_name, **self.init_kwargs)
        swanlab.config["FRAMEWORK"] = "accelerate"
        return run

    def _config(self, values):
        import swanlab

        swanlab.config.update(values, allow_val_change=True)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        import swanlab

        for tag, images in values.items():
            self._run.log({tag: [swanlab.Image(img) for img in images]}, step=step, **kwargs)


class CometMLTracker(_BackendTracker):
    name = "comet_ml"

    def _open(self):
        import comet_ml

        opener = getattr(comet_ml, "start", None)
        if opener is not None:
            return opener(project_name=self.run_name, **self.init_kwargs)
        return comet_ml.Experiment(project_name=self.run_name, **self.init_kwargs)

    def _config(self, values):
        self._run.log_parameters(values)

    def _scalars(self, values, step, **kwargs):
        exp = self._run
        if step is not None:
            exp.set_step(step)
        for key, value in _scalarize(values).items():
            if isinstance(value, (int, float)):
                exp.log_metric(key, value, step=step, **kwargs)
            elif isinstance(value, str):
                exp.log_other(key, value, **kwargs)
            elif isinstance(value, dict):
                exp.log_metrics(value, step=step, **kwargs)

    def _close(self):
        self._run.end()


class AimTracker(_BackendTracker):
    name = "aim"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Optional[Union[str, os.PathLike]] = ".", **kwargs):
        super().__init__(run_name, logging_dir, **kwargs)

    def _open(self):
        from aim import Run

        run = Run(repo=self.logging_dir, **self.init_kwargs)
        run.name = self.run_name
        return run

    def _config(self, values):
        self._run["hparams"] = values

    def _scalars(self, values, step, **kwargs):
        for key, value in values.items():
            self._run.track(value, name=key, step=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, kwargs: Optional[dict] = None):
        """``values``: name → image or ``(image, caption)``; ``kwargs`` may hold ``{"aim_image": {...}, "track": {...}}``."""
        import aim

        extra = {"aim_image": {}, "track": {}}
        extra.update(kwargs or {})
        for key, value in values.items():
            image, caption = value if isinstance(value, tuple) else (value, "")
            self._run.track(aim.Image(image, caption=caption, **extra["aim_image"]), name=key, step=step, **extra["track"])

    def _close(self):
        self._run.close()


class MLflowTracker(_BackendTracker):
    name = "mlflow"
    _PARAMS_PER_CALL = 100  # MLflow's batch limit for log_params

    def __init__(self, experiment_name: str = None, logging_dir: Optional[Union[str, os.PathLike]] = None, run_id: Optional[str] = None,
                 tags: Optional[Union[dict, str]] = None, nested_run: Optional[bool] = False, run_name: Optional[str] = None,
                 description: Optional[str] = None):
        env = os.environ
        experiment_name = env.get("MLFLOW_EXPERIMENT_NAME", experiment_name)
        run_id = env.get("MLFLOW_RUN_ID", run_id)
        tags = env.get("MLFLOW_TAGS", tags)
        if isinstance(tags, str):
            tags = json.loads(tags)
        nested_run = env.get("MLFLOW_NESTED_RUN", nested_run)
        super().__init__(run_name, logging_dir)
        self.experiment_name, self.run_id, self.tags, self.nested_run, self.description = experiment_name, run_id, tags, nested_run, description

    def _open(self):
        import mlflow

        hits = mlflow.search_experiments(filter_string=f"name = '{self.experiment_name}'")
        if hits:
            if len(hits) > 1:
                raise ValueError("There seems to be multiple experiments with the same name")
            experiment_id = hits[0].experiment_id
        else:
            experiment_id = mlflow.create_experiment(name=self.experiment_name, artifact_location=self.logging_dir, tags=self.tags)
        return mlflow.start_run(run_id=self.run_id, experiment_id=experiment_id, run_name=self.run_name, nested=self.nested_run,
                                tags=self.tags, description=self.description)

    def _config(self, values):
        import mlflow

        limit = mlflow.utils.validation.MAX_PARAM_VAL_LENGTH
        kept = {}
        for key, value in values.items():
            if len(str(value)) > limit:
                logger.warning_once(f'Accelerate is attempting to log a value of "{value}" for key "{key}" as a parameter. MLflow\'s '
                                    f"log_param() only accepts values no longer than {limit} characters so we dropped this attribute.")
            else:
                kept[key] = value
        items = list(kept.items())
        for i in range(0, len(items), self._PARAMS_PER_CALL):
            mlflow.log_params(dict(items[i : i + self._PARAMS_PER_CALL]))

    def _scalars(self, values, step, **kwargs):
        import mlflow

        numbers = {}
        for key, value in _scalarize(values).items():
            if isinstance(value, (int, float)):
                numbers[key] = value
            else:
                logger.warning_once(f'MLflowTracker is attempting to log a value of "{value}" of type {type(value)} for key "{key}" as a '
                                    "metric. MLflow's log_metric() only accepts float and int types so we dropped this attribute.")
        mlflow.log_metrics(numbers, step=step)

    @on_main_process
    def log_figure(self, figure: Any, artifact_file: str, **save_kwargs):
        import mlflow

        mlflow.log_figure(figure, artifact_file, **save_kwargs)

    @on_main_process
    def log_artifacts(self, local_dir: str, artifact_path: Optional[str] = None):
        import mlflow

        mlflow.log_artifacts(local_dir, artifact_path=artifact_path)

    @on_main_process
    def log_artifact(self, local_path: str, artifact_path: Optional[str] = None):
        import mlflow

        mlflow.log_artifact(local_path, artifact_path=artifact_path)

    def _close(self):
        import mlflow

        mlflow.end_run()


class ClearMLTracker(_BackendTracker):
    name = "clearml"
    _SPLITS = ("eval", "test", "train")

    def _open(self):
        from clearml import Task

        self._created_here = False
        task = Task.current_task()
        if task is None:
            self._created_here = True
            opts = dict(self.init_kwargs)
            opts.setdefault("project_name", os.environ.get("CLEARML_PROJECT", self.run_name))
            opts.setdefault("task_name", os.environ.get("CLEARML_TASK", self.run_name))
            task = Task.init(**opts)
        return task

    def _config(self, values):
        return self._run.connect_configuration(values)

    @classmethod
    def _title_series(cls, key: str):
        """``eval_loss`` → title ``loss``, series ``eval`` (so train / eval curves share one plot)."""
        for split in cls._SPLITS:
            if key.startswith(split + "_"):
                return key[len(split) + 1 :], split
        return key, "train"

    def _scalars(self, values, step, **kwargs):
        sink = self._run.get_logger()
        for key, value in _scalarize(values).items():
            if not isinstance(value, (int, float)):
                logger.warning_once(f'Accelerator is attempting to log a value of "{value}" of type {type(value)} for key "{key}" as a '
                                    "scalar. This invocation of ClearML logger's  report_scalar() is incorrect so we dropped this attribute.")
            elif step is None:
                sink.report_single_value(name=key, value=value, **kwargs)
            else:
                title, series = self._title_series(key)
                sink.report_scalar(title=title, series=series, value=value, iteration=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        sink = self._run.get_logger()
        for key, image in values.items():
            title, series = self._title_series(key)
            sink.report_image(title=title, series=series, iteration=step, image=image, **kwargs)

    @on_main_process
    def log_table(self, table_name: str, columns=None, data=None, dataframe=None, step: Optional[int] = None, **kwargs):
        if dataframe is None:
            if data is None:
                raise ValueError("`ClearMLTracker.log_table` requires that `data` to be supplied if `dataframe` is `None`")
            dataframe = [columns] + data if columns else data
        title, series = self._title_series(table_name)
        self._run.get_logger().report_table(title=title, series=series, table_plot=dataframe, iteration=step, **kwargs)

    def _close(self):
        if self._run is not None and self._created_here:
            self._run.close()


class DVCLiveTracker(_BackendTracker):
    name = "dvclive"

    def __init__(self, run_name: Optional[str] = None, live: Optional[Any] = None, **kwargs):
        super().__init__(run_name, None, **kwargs)
        self._given = live

    def _open(self):
        if self._given is not None:
            return self._given
        from dvclive import Live

        return Live(**self.init_kwargs)

    def _config(self, values):
        self._run.log_params(values)

    def _scalars(self, values, step, **kwargs):
        live = self._run
        if step is not None:
            live.step = step
        for key, value in _scalarize(values).items():
            if isinstance(value, (int, float, str)):
                live.log_metric(key, value, **kwargs)
            else:
                logger.warning_once(f'Accelerator attempted to log a value of "{value}" of type {type(value)} for key "{key}" as a scalar. '
                                    "This invocation of DVCLive's Live.log_metric() is incorrect so we dropped this attribute.")
        live.next_step()

    def _close(self):
        self._run.end()


# =========================================================================== registry
LOGGER_TYPE_TO_CLASS = {
    "aim": AimTracker,
    "comet_ml": CometMLTracker,
    "mlflow": MLflowTracker,
    "tensorboard": TensorBoardTracker,
    "wandb": WandBTracker,
    "clearml": ClearMLTracker,
    "dvclive": DVCLiveTracker,
    "swanlab": SwanLabTracker,
    "trackio": TrackioTracker,
    "jsonl": JSONLTracker,
}

_PROBES = {
    LoggerType.TENSORBOARD: "is_tensorboard_available", LoggerType.WANDB: "is_wandb_available", LoggerType.COMETML: "is_comet_ml_available",
    LoggerType.AIM: "is_aim_available", LoggerType.MLFLOW: "is_mlflow_available", LoggerType.CLEARML: "is_clearml_available",
    LoggerType.DVCLIVE: "is_dvclive_available", LoggerType.SWANLAB: "is_swanlab_available", LoggerType.TRACKIO: "is_trackio_available",
}


def _probe_installed() -> list:
    found = []
    for kind, probe in _PROBES.items():
        try:
            if getattr(_imports, probe)():
                found.append(kind)
        except Exception:  # noqa: BLE001 - a broken optional package must not break tracking as a whole
            pass
    found.append(LoggerType.JSONL)
    return found


_available_trackers = _probe_installed()


def get_available_trackers() -> list:
    """``LoggerType`` s whose package is importable here (JSONL always is)."""
    return _available_trackers


def filter_trackers(log_with, logging_dir: Union[str, os.PathLike] = None) -> list:
    """Normalise ``Accelerator(log_with=...)``: ``"all"`` → every installed integration (+ any tracker instances passed along);
    names / ``LoggerType`` s are kept when their package is installed (a missing one is skipped with a debug message) and
    rejected when they need a ``logging_dir`` that was not given; ``GeneralTracker`` instances pass through."""
    if log_with is None:
        return []
    wanted = list(log_with) if isinstance(log_with, (list, tuple)) else [log_with]
    is_instance = lambda item: isinstance(item, GeneralTracker)  # noqa: E731
    if "all" in wanted or LoggerType.ALL in wanted:
        return [item for item in wanted if is_instance(item)] + get_available_trackers()
    chosen: list = []
    for item in wanted:
        if is_instance(item):
            chosen.append(item)
            continue
        if item not in LoggerType:
            raise ValueError(f"Unsupported logging capability: {item}. Choose between {LoggerType.list()}")
        kind = LoggerType(item)
        if kind in chosen:
            continue
        if kind not in get_available_trackers():
            logger.debug(f"Tried adding logger {kind}, but package is unavailable in the system.")
            continue
        if LOGGER_TYPE_TO_CLASS[str(kind)].requires_logging_directory and logging_dir is None:
            raise ValueError(f"Logging with `{kind}` requires a `logging_dir` to be passed in.")
        chosen.append(kind)
    return chosen
