from .args import EvalPrediction, TrainerState, TrainingArguments, TrainOutput  # noqa: F401
from .trainer import LambdaSchedule, Trainer  # noqa: F401
