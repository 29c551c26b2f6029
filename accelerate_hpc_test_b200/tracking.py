"""Experiment trackers behind ``Accelerator.init_trackers / log / end_training``.

Contract parity with reference src/accelerate/tracking.py: the ``GeneralTracker`` ABC :101-179 (class
attributes ``name``, ``requires_logging_directory``, ``main_process_only``; methods ``start``,
``store_init_configuration``, ``log``, ``finish``; property ``tracker``), the ``on_main_process``
decorator :77-93, the built-in integrations :182-1246, ``LOGGER_TYPE_TO_CLASS`` :1249-1259 and
``filter_trackers`` :1262-1317.  Added: ``JSONLTracker`` — dependency-free, one JSON object per
``log`` call, also used for the built-in perf counters (tokens/s, step ms, exposed-comm ms).
"""

from __future__ import annotations

import json
import os
import time
from functools import wraps
from typing import Any, Optional, Union

from .logging import get_logger
from .state import PartialState
from .utils.dataclasses import LoggerType
from .utils.imports import (
    is_aim_available,
    is_clearml_available,
    is_comet_ml_available,
    is_dvclive_available,
    is_mlflow_available,
    is_swanlab_available,
    is_tensorboard_available,
    is_trackio_available,
    is_wandb_available,
)
from .utils.operations import listify

logger = get_logger(__name__)

_available_trackers = []
for _flag, _kind in (
    (is_tensorboard_available, LoggerType.TENSORBOARD), (is_wandb_available, LoggerType.WANDB),
    (is_comet_ml_available, LoggerType.COMETML), (is_aim_available, LoggerType.AIM), (is_mlflow_available, LoggerType.MLFLOW),
    (is_clearml_available, LoggerType.CLEARML), (is_dvclive_available, LoggerType.DVCLIVE), (is_swanlab_available, LoggerType.SWANLAB),
    (is_trackio_available, LoggerType.TRACKIO),
):
    try:
        if _flag():
            _available_trackers.append(_kind)
    except Exception:  # pragma: no cover
        pass
_available_trackers.append(LoggerType.JSONL)


def on_main_process(function):
    """Run the tracker method only on the main process when the tracker says ``main_process_only``."""

    @wraps(function)
    def execute_on_main_process(self, *args, **kwargs):
        if getattr(self, "main_process_only", False):
            return PartialState().on_main_process(function)(self, *args, **kwargs)
        return function(self, *args, **kwargs)

    return execute_on_main_process


def get_available_trackers():
    return _available_trackers


class GeneralTracker:
    """Base class of every tracker.  Subclasses set ``name``, ``requires_logging_directory`` and
    implement ``store_init_configuration`` and ``log`` (``start`` / ``finish`` are optional)."""

    main_process_only = True

    def __init__(self, _blank: bool = False):
        if not _blank:
            err = ""
            if not hasattr(self, "name"):
                err += "`name`"
            if not hasattr(self, "requires_logging_directory"):
                if len(err) > 0:
                    err += ", "
                err += "`requires_logging_directory`"
            if "tracker" not in dir(self):
                if len(err) > 0:
                    err += ", "
                err += "`tracker`"
            if len(err) > 0:
                raise NotImplementedError(f"The implementation for this tracker class is missing the following required attributes. Please define them in the class definition: {err}")

    def start(self):
        pass

    def store_init_configuration(self, values: dict):
        pass

    def log(self, values: dict, step: Optional[int], **kwargs):
        pass

    def finish(self):
        pass


class JSONLTracker(GeneralTracker):
    """Appends one JSON line per ``log`` call to ``{logging_dir}/{run_name}/metrics.jsonl``."""

    name = "jsonl"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Union[str, os.PathLike], **kwargs):
        super().__init__()
        self.run_name = run_name
        self.logging_dir = os.path.join(str(logging_dir), run_name)
        self._fh = None

    @on_main_process
    def start(self):
        os.makedirs(self.logging_dir, exist_ok=True)
        self._fh = open(os.path.join(self.logging_dir, "metrics.jsonl"), "a", buffering=1)

    @property
    def tracker(self):
        return self._fh

    @on_main_process
    def store_init_configuration(self, values: dict):
        with open(os.path.join(self.logging_dir, "hparams.json"), "w") as f:
            json.dump(_plain(values), f, indent=1, default=str)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        if self._fh is None:
            self.start()
        rec = {"step": step, "time": time.time()}
        rec.update(_plain(values))
        self._fh.write(json.dumps(rec, default=str) + "\n")

    @on_main_process
    def finish(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None


def _plain(values):
    out = {}
    for k, v in values.items():
        if hasattr(v, "item") and getattr(v, "numel", lambda: 2)() == 1:
            v = v.item()
        elif hasattr(v, "tolist"):
            v = v.tolist()
        out[k] = v
    return out


class TensorBoardTracker(GeneralTracker):
    name = "tensorboard"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Union[str, os.PathLike], **kwargs):
        super().__init__()
        self.run_name = run_name
        self.logging_dir_param = logging_dir
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        try:
            from torch.utils import tensorboard
        except ModuleNotFoundError:
            import tensorboardX as tensorboard
        self.logging_dir = os.path.join(self.logging_dir_param, self.run_name)
        self.writer = tensorboard.SummaryWriter(self.logging_dir, **self.init_kwargs)
        logger.debug(f"Initialized TensorBoard project {self.run_name} logging to {self.logging_dir}")

    @property
    def tracker(self):
        return self.writer

    @on_main_process
    def store_init_configuration(self, values: dict):
        self.writer.add_hparams(values, metric_dict={})
        self.writer.flush()
        project_run_name = time.time()
        dir_name = os.path.join(self.logging_dir, str(project_run_name))
        os.makedirs(dir_name, exist_ok=True)
        import yaml

        with open(os.path.join(dir_name, "hparams.yml"), "w") as outfile:
            try:
                yaml.dump(values, outfile)
            except yaml.representer.RepresenterError:
                logger.error("Serialization to store hyperparameters failed")
                raise

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        values = listify(values)
        for k, v in values.items():
            if isinstance(v, (int, float)):
                self.writer.add_scalar(k, v, global_step=step, **kwargs)
            elif isinstance(v, str):
                self.writer.add_text(k, v, global_step=step, **kwargs)
            elif isinstance(v, dict):
                self.writer.add_scalars(k, v, global_step=step, **kwargs)
        self.writer.flush()

    @on_main_process
    def log_images(self, values: dict, step: Optional[int], **kwargs):
        for k, v in values.items():
            self.writer.add_images(k, v, global_step=step, **kwargs)

    @on_main_process
    def finish(self):
        self.writer.close()


class WandBTracker(GeneralTracker):
    name = "wandb"
    requires_logging_directory = False
    main_process_only = True

    def __init__(self, run_name: str, **kwargs):
        super().__init__()
        self.run_name = run_name
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        import wandb

        self.run = wandb.init(project=self.run_name, **self.init_kwargs)

    @property
    def tracker(self):
        return self.run

    @on_main_process
    def store_init_configuration(self, values: dict):
        import wandb

        wandb.config.update(values, allow_val_change=True) if hasattr(wandb, "config") else self.run.config.update(values, allow_val_change=True)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        self.run.log(values, step=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        import wandb

        for k, v in values.items():
            self.log({k: [wandb.Image(image) for image in v]}, step=step, **kwargs)

    @on_main_process
    def log_table(self, table_name: str, columns=None, data=None, dataframe=None, step: Optional[int] = None, **kwargs):
        import wandb

        self.log({table_name: wandb.Table(columns=columns, data=data, dataframe=dataframe)}, step=step, **kwargs)

    @on_main_process
    def finish(self):
        self.run.finish()


class TrackioTracker(GeneralTracker):
    name = "trackio"
    requires_logging_directory = False

    def __init__(self, run_name: str, **kwargs):
        super().__init__()
        self.run_name = run_name
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        import trackio

        self.run = trackio.init(project=self.run_name, **self.init_kwargs)

    @property
    def tracker(self):
        return self.run

    @on_main_process
    def store_init_configuration(self, values: dict):
        import trackio

        trackio.config.update(values, allow_val_change=True)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        self.run.log(values, **kwargs)

    @on_main_process
    def finish(self):
        self.run.finish()


class CometMLTracker(GeneralTracker):
    name = "comet_ml"
    requires_logging_directory = False

    def __init__(self, run_name: str, **kwargs):
        super().__init__()
        self.run_name = run_name
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        import comet_ml

        self.writer = comet_ml.start(project_name=self.run_name, **self.init_kwargs) if hasattr(comet_ml, "start") else comet_ml.Experiment(project_name=self.run_name, **self.init_kwargs)

    @property
    def tracker(self):
        return self.writer

    @on_main_process
    def store_init_configuration(self, values: dict):
        self.writer.log_parameters(values)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        if step is not None:
            self.writer.set_step(step)
        for k, v in values.items():
            if isinstance(v, (int, float)):
                self.writer.log_metric(k, v, step=step, **kwargs)
            elif isinstance(v, str):
                self.writer.log_other(k, v, **kwargs)
            elif isinstance(v, dict):
                self.writer.log_metrics(v, step=step, **kwargs)

    @on_main_process
    def finish(self):
        self.writer.end()


class AimTracker(GeneralTracker):
    name = "aim"
    requires_logging_directory = True

    def __init__(self, run_name: str, logging_dir: Optional[Union[str, os.PathLike]] = ".", **kwargs):
        super().__init__()
        self.run_name = run_name
        self.aim_repo_path = logging_dir
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        from aim import Run

        self.writer = Run(repo=self.aim_repo_path, **self.init_kwargs)
        self.writer.name = self.run_name

    @property
    def tracker(self):
        return self.writer

    @on_main_process
    def store_init_configuration(self, values: dict):
        self.writer["hparams"] = values

    @on_main_process
    def log(self, values: dict, step: Optional[int], **kwargs):
        for key, value in values.items():
            self.writer.track(value, name=key, step=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, kwargs: Optional[dict] = None):
        import aim

        aim_image_kw, track_kw = {}, {}
        if kwargs is not None:
            aim_image_kw = kwargs.get("aim_image", {})
            track_kw = kwargs.get("track", {})
        for key, value in values.items():
            if isinstance(value, tuple):
                img, caption = value
            else:
                img, caption = value, ""
            self.writer.track(aim.Image(img, caption=caption, **aim_image_kw), name=key, step=step, **track_kw)

    @on_main_process
    def finish(self):
        self.writer.close()


class MLflowTracker(GeneralTracker):
    name = "mlflow"
    requires_logging_directory = False

    def __init__(self, experiment_name: str = None, logging_dir: Optional[Union[str, os.PathLike]] = None, run_id: Optional[str] = None,
                 tags: Optional[Union[dict, str]] = None, nested_run: Optional[bool] = False, run_name: Optional[str] = None,
                 description: Optional[str] = None):
        super().__init__()
        experiment_name = os.environ.get("MLFLOW_EXPERIMENT_NAME", experiment_name)
        run_id = os.environ.get("MLFLOW_RUN_ID", run_id)
        tags = os.environ.get("MLFLOW_TAGS", tags)
        if isinstance(tags, str):
            tags = json.loads(tags)
        nested_run = os.environ.get("MLFLOW_NESTED_RUN", nested_run)
        self.experiment_name = experiment_name
        self.logging_dir = logging_dir
        self.run_id = run_id
        self.tags = tags
        self.nested_run = nested_run
        self.run_name = run_name
        self.description = description

    @on_main_process
    def start(self):
        import mlflow

        exps = mlflow.search_experiments(filter_string=f"name = '{self.experiment_name}'")
        if len(exps) > 0:
            if len(exps) > 1:
                logger.warning("Multiple experiments with the same name found. Using first one.")
            experiment_id = exps[0].experiment_id
        else:
            experiment_id = mlflow.create_experiment(name=self.experiment_name, artifact_location=self.logging_dir, tags=self.tags)
        self.active_run = mlflow.start_run(run_id=self.run_id, experiment_id=experiment_id, run_name=self.run_name,
                                           nested=self.nested_run, tags=self.tags, description=self.description)

    @property
    def tracker(self):
        return self.active_run

    @on_main_process
    def store_init_configuration(self, values: dict):
        import mlflow

        for name, value in list(values.items()):
            if len(str(value)) > mlflow.utils.validation.MAX_PARAM_VAL_LENGTH:
                logger.warning_once(f'Accelerate is attempting to log a value of "{value}" for key "{name}" as a parameter. MLflow\'s log_param() only accepts values no longer than {mlflow.utils.validation.MAX_PARAM_VAL_LENGTH} characters so we dropped this attribute.')
                del values[name]
        values_list = list(values.items())
        for i in range(0, len(values_list), mlflow.utils.validation.MAX_PARAMS_TAGS_PER_BATCH):
            mlflow.log_params(dict(values_list[i : i + mlflow.utils.validation.MAX_PARAMS_TAGS_PER_BATCH]))

    @on_main_process
    def log(self, values: dict, step: Optional[int]):
        import mlflow

        metrics = {}
        for k, v in values.items():
            if isinstance(v, (int, float)):
                metrics[k] = v
            else:
                logger.warning_once(f'MLflowTracker is attempting to log a value of "{v}" of type {type(v)} for key "{k}" as a metric. MLflow\'s log_metric() only accepts float and int types so we dropped this attribute.')
        mlflow.log_metrics(metrics, step=step)

    @on_main_process
    def log_figure(self, figure: Any, artifact_file: str, **save_kwargs):
        import mlflow

        mlflow.log_figure(figure=figure, artifact_file=artifact_file, **save_kwargs)

    @on_main_process
    def log_artifacts(self, local_dir: str, artifact_path: Optional[str] = None):
        import mlflow

        mlflow.log_artifacts(local_dir=local_dir, artifact_path=artifact_path)

    @on_main_process
    def log_artifact(self, local_path: str, artifact_path: Optional[str] = None):
        import mlflow

        mlflow.log_artifact(local_path=local_path, artifact_path=artifact_path)

    @on_main_process
    def finish(self):
        import mlflow

        mlflow.end_run()


class ClearMLTracker(GeneralTracker):
    name = "clearml"
    requires_logging_directory = False

    def __init__(self, run_name: str = None, **kwargs):
        super().__init__()
        self.user_provided_run_name = run_name
        self._initialized_externally = False
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        from clearml import Task

        current_task = Task.current_task()
        if current_task:
            self._initialized_externally = True
            self.task = current_task
            return
        task_init_args = {**self.init_kwargs}
        task_init_args.setdefault("project_name", os.environ.get("CLEARML_PROJECT", self.user_provided_run_name))
        task_init_args.setdefault("task_name", os.environ.get("CLEARML_TASK", self.user_provided_run_name))
        self.task = Task.init(**task_init_args)

    @property
    def tracker(self):
        return self.task

    @on_main_process
    def store_init_configuration(self, values: dict):
        return self.task.connect_configuration(values)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        clearml_logger = self.task.get_logger()
        for k, v in values.items():
            if not isinstance(v, (int, float)):
                logger.warning_once(f"Accelerator is attempting to log a value of \"{v}\" of type {type(v)} for key \"{k}\" as a scalar. This invocation of ClearML logger's  report_scalar() is incorrect so we dropped this attribute.")
                continue
            if step is None:
                clearml_logger.report_single_value(name=k, value=v, **kwargs)
                continue
            title, series = ClearMLTracker._get_title_series(k)
            clearml_logger.report_scalar(title=title, series=series, value=v, iteration=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        clearml_logger = self.task.get_logger()
        for k, v in values.items():
            title, series = ClearMLTracker._get_title_series(k)
            clearml_logger.report_image(title=title, series=series, iteration=step, image=v, **kwargs)

    @on_main_process
    def log_table(self, table_name: str, columns=None, data=None, dataframe=None, step: Optional[int] = None, **kwargs):
        to_report = dataframe
        if dataframe is None:
            if data is None:
                raise ValueError("`ClearMLTracker.log_table` requires that `data` to be supplied if `dataframe` is `None`")
            to_report = [columns] + data if columns else data
        title, series = ClearMLTracker._get_title_series(table_name)
        self.task.get_logger().report_table(title=title, series=series, table_plot=to_report, iteration=step, **kwargs)

    @on_main_process
    def finish(self):
        if self.task and not self._initialized_externally:
            self.task.close()

    @staticmethod
    def _get_title_series(name):
        for prefix in ["eval", "test", "train"]:
            if name.startswith(prefix + "_"):
                return name[len(prefix) + 1 :], prefix
        return name, "train"


class DVCLiveTracker(GeneralTracker):
    name = "dvclive"
    requires_logging_directory = False

    def __init__(self, run_name: Optional[str] = None, live: Optional[Any] = None, **kwargs):
        super().__init__()
        self.live = live
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        from dvclive import Live

        self.live = self.live if self.live is not None else Live(**self.init_kwargs)

    @property
    def tracker(self):
        return self.live

    @on_main_process
    def store_init_configuration(self, values: dict):
        self.live.log_params(values)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        from dvclive.plots import Metric

        if step is not None:
            self.live.step = step
        for k, v in values.items():
            if Metric.could_log(v):
                self.live.log_metric(k, v, **kwargs)
            else:
                logger.warning_once(f'Accelerator attempted to log a value of "{v}" of type {type(v)} for key "{k}" as a scalar. This invocation of DVCLive\'s Live.log_metric() is incorrect so we dropped this attribute.')
        self.live.next_step()

    @on_main_process
    def finish(self):
        self.live.end()


class SwanLabTracker(GeneralTracker):
    name = "swanlab"
    requires_logging_directory = False

    def __init__(self, run_name: str, **kwargs):
        super().__init__()
        self.run_name = run_name
        self.init_kwargs = kwargs

    @on_main_process
    def start(self):
        import swanlab

        self.run = swanlab.init(project=self.run_name, **self.init_kwargs)
        swanlab.config["FRAMEWORK"] = "accelerate"

    @property
    def tracker(self):
        return self.run

    @on_main_process
    def store_init_configuration(self, values: dict):
        import swanlab

        swanlab.config.update(values, allow_val_change=True)

    @on_main_process
    def log(self, values: dict, step: Optional[int] = None, **kwargs):
        self.run.log(values, step=step, **kwargs)

    @on_main_process
    def log_images(self, values: dict, step: Optional[int] = None, **kwargs):
        import swanlab

        for k, v in values.items():
            self.log({k: [swanlab.Image(image) for image in v]}, step=step, **kwargs)

    @on_main_process
    def finish(self):
        self.run.finish()


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


def filter_trackers(log_with: list, logging_dir: Union[str, os.PathLike] = None):
    """Resolve ``"all"`` / names / ``LoggerType`` s / tracker instances to the usable ones."""
    loggers = []
    if log_with is not None:
        if not isinstance(log_with, (list, tuple)):
            log_with = [log_with]
        if "all" in log_with or LoggerType.ALL in log_with:
            loggers = [o for o in log_with if issubclass(type(o), GeneralTracker)] + get_available_trackers()
        else:
            for log_type in log_with:
                if log_type not in LoggerType and not issubclass(type(log_type), GeneralTracker):
                    raise ValueError(f"Unsupported logging capability: {log_type}. Choose between {LoggerType.list()}")
                if issubclass(type(log_type), GeneralTracker):
                    loggers.append(log_type)
                else:
                    log_type = LoggerType(log_type)
                    if log_type not in loggers:
                        if log_type in get_available_trackers():
                            tracker_init = LOGGER_TYPE_TO_CLASS[str(log_type)]
                            if tracker_init.requires_logging_directory:
                                if logging_dir is None:
                                    raise ValueError(f"Logging with `{log_type}` requires a `logging_dir` to be passed in.")
                            loggers.append(log_type)
                        else:
                            logger.debug(f"Tried adding logger {log_type}, but package is unavailable in the system.")
    return loggers
