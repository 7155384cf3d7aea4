from .averager import AveragingMode, DecentralizedAverager
from .compression import (CompressionBase, Float16Compression, NoCompression, SizeAdaptiveCompression,
                          Uniform8BitQuantization)
from .optimizer import (CollaborativeOptimizer, GradientAverager, HivemindGradScaler, Optimizer,
                        TrainingStateAverager)
from .progress_tracker import LocalTrainingProgress, PerformanceEMA, ProgressTracker

__all__ = ["Optimizer", "CollaborativeOptimizer", "GradientAverager", "TrainingStateAverager", "HivemindGradScaler",
           "ProgressTracker", "PerformanceEMA", "LocalTrainingProgress", "DecentralizedAverager", "AveragingMode",
           "CompressionBase", "NoCompression", "Float16Compression", "Uniform8BitQuantization",
           "SizeAdaptiveCompression"]
