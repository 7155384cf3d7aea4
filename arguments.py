"""Command-line configuration: the flag surface of the reference (reference arguments.py:8-165).

Every flag name and default is kept, so launch lines written for the reference keep working.  Differences:
``HFTrainerArguments`` no longer inherits from ``transformers.TrainingArguments`` (importing that class needs
``accelerate``, and the reference only ever reads a handful of its fields - they are declared here explicitly);
``--model_preset`` / ``--micro_batches_in_flight`` style additions are optional and default to reference behaviour.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass, field
from typing import List, Optional

import torch


def _opt(default, help_text: str):
    if isinstance(default, (list, dict)):
        return field(default_factory=lambda: type(default)(default), metadata={"help": help_text})
    return field(default=default, metadata={"help": help_text})


@dataclass
class TrainingArgumentsBase:
    """The subset of ``transformers.TrainingArguments`` the reference reads (task.py:53-56,88; run_trainer.py:34,37;
    utils.py:59-63)."""

    seed: int = _opt(42, "Random seed used for parameter initialisation")
    local_rank: int = _opt(-1, "Rank inside a multi-process launch (-1: not distributed)")
    run_name: Optional[str] = _opt(None, "Run name for experiment trackers")
    do_eval: bool = _opt(False, "Evaluation is not supported by the collaborative trainer")
    no_cuda: bool = _opt(False, "Run on CPU even when a GPU is visible")
    bf16: bool = _opt(True, "bf16 GEMM operands with fp32 master weights (the B200 compute dtype)")

    @property
    def device(self) -> torch.device:
        if self.no_cuda or not torch.cuda.is_available():
            return torch.device("cpu")
        return torch.device("cuda", max(self.local_rank, 0))

    @property
    def n_gpu(self) -> int:
        return 0 if self.no_cuda else torch.cuda.device_count()


@dataclass
class HFTrainerArguments(TrainingArgumentsBase):
    """Local training hyper-parameters."""

    dataloader_num_workers: int = 1
    per_device_train_batch_size: int = 2
    per_device_eval_batch_size: int = 2
    gradient_accumulation_steps: int = 1
    text_seq_length: int = 256

    # optimisation of the DALL-E model
    learning_rate: float = 0.0025
    adam_beta1: float = 0.9
    adam_beta2: float = 0.96
    max_grad_norm: float = 4.0
    weight_decay: float = 0.045
    total_steps: int = _opt(31250, "Number of collaborative optimizer updates; defines the learning-rate schedule")
    warmup_steps: int = 3125
    adam_epsilon: float = 1e-6
    clamp_value: float = 10000.0

    fp16: bool = False
    do_train: bool = True
    logging_steps: int = 100
    max_steps: int = 10 ** 20
    save_steps: int = 10 ** 20
    save_total_limit: int = 2
    output_dir: str = "outputs"

    # additions (defaults reproduce the reference)
    model_preset: str = _opt("ttt", "Model architecture preset: ttt (reference model) | tiny | small")
    synthetic_data: bool = _opt(False, "Train on synthetic token ids / VQGAN codes instead of the LAION stream")

    @property
    def batch_size_per_step(self) -> int:
        """Samples this peer contributes per optimizer.step() call (all visible GPUs count, as in the reference)."""
        per_step = self.per_device_train_batch_size * self.gradient_accumulation_steps
        return per_step * max(1, 0 if self.no_cuda else torch.cuda.device_count())


@dataclass
class TPUTrainerArguments(HFTrainerArguments):
    num_tpus: int = _opt(8, "Accelerator workers fed by one host process (TPU cores in the reference; GPUs here)")
    wandb_project: str = "huggingface"

    @property
    def batch_size_per_step(self) -> int:
        return self.per_device_train_batch_size * self.gradient_accumulation_steps * self.num_tpus


@dataclass
class CollaborativeArguments:
    """Keyword arguments forwarded verbatim to the collaborative optimizer."""

    target_batch_size: int = _opt(4096, "Perform optimizer step after all peers collectively accumulate this many samples")
    matchmaking_time: float = _opt(15.0, "Averaging group will wait for stragglers for at most this many seconds")
    allreduce_timeout: float = _opt(60, "Give up on a given all-reduce round after this many seconds")
    averaging_timeout: float = _opt(180, "Give up on averaging step after this many seconds")
    reuse_grad_buffers: bool = _opt(True, "Accumulate gradients across local steps directly in the model's .grad "
                                          "buffers (saves memory; incompatible with custom loss scalers)")


@dataclass
class BasePeerArguments:
    """Arguments shared by trainers and auxiliary peers."""

    experiment_prefix: str = _opt("my-model", "A unique experiment name, used as prefix for all DHT keys")
    tokenizer_path: Optional[str] = _opt("t5-small", "Path to the tokenizer")
    cache_dir: Optional[str] = _opt("./cache", "Path to the cache")
    authorize: bool = _opt(True, "Whether or not to use HF authorizer")
    client_mode: bool = _opt(False, "If True, runs training without incoming connections, in a firewall-compatible mode")
    initial_peers: List[str] = _opt([], "Multiaddrs of the peers that will welcome you into the existing "
                                        "collaboration. Example: /ip4/203.0.113.1/tcp/31337/p2p/XXXX")
    use_ipfs: bool = _opt(False, "Use IPFS to find initial_peers (needs internet; rejected on isolated hosts)")
    host_maddrs: List[str] = _opt(["/ip4/0.0.0.0/tcp/0"], "Multiaddrs to listen for external connections")
    announce_maddrs: List[str] = _opt([], "Visible multiaddrs the host announces for external connections")
    identity_path: Optional[str] = _opt(None, "Path to a pre-generated private key file; makes the peer ID deterministic")


@dataclass
class TrainingPeerArguments(BasePeerArguments):
    statistics_expiration: float = _opt(600, "Statistics will be removed if not updated in this many seconds")
    backup_every_steps: Optional[int] = _opt(None, "Update training state backup on disk once in this many global "
                                                   "steps (default = do not update local state)")
    state_path: str = _opt("state.zip", "Load this state upon init and when recovering from NaN parameters")


@dataclass
class AuxiliaryPeerArguments(BasePeerArguments):
    """run_aux_peer.py: connects peers, tracks learning curves, assists all-reduce, uploads checkpoints."""

    refresh_period: float = _opt(10, "Period (in seconds) for fetching the keys from DHT")
    wandb_project: Optional[str] = _opt(None, "Name of Weights & Biases project to report the training progress to")
    save_checkpoint_step_interval: int = _opt(2, "Frequency (in steps) of fetching and saving state from peers")
    repo_url: Optional[str] = _opt(None, "URL of Hugging Face Hub repository to upload the model and optimizer states")
    local_path: Optional[str] = _opt("Repo", "Path to local repository to store the model and optimizer states")
    upload_interval: Optional[float] = _opt(None, "Frequency (in seconds) of uploading the model to Hub")
    store_checkpoints: bool = _opt(True, "If True, enables CheckpointHandler")
    assist_in_averaging: bool = _opt(False, "If True, this peer will facilitate averaging for other (training) peers")
    assist_refresh: float = _opt(1.0, "Period (in seconds) for tryin to assist averaging")


def parse_dataclasses(classes, args=None):
    """``HfArgumentParser`` when transformers is importable, a small argparse equivalent otherwise."""
    try:
        from transformers import HfArgumentParser

        return HfArgumentParser(classes).parse_args_into_dataclasses(args=args)
    except ImportError:  # pragma: no cover
        import argparse

        parser = argparse.ArgumentParser()
        for cls in classes:
            for f in dataclasses.fields(cls):
                default = f.default if f.default is not dataclasses.MISSING else f.default_factory()
                kw = {"help": f.metadata.get("help")}
                if isinstance(default, bool):
                    kw.update(type=lambda s: str(s).lower() in ("1", "true", "yes"), nargs="?", const=True)
                elif isinstance(default, list):
                    kw.update(nargs="*", type=str)
                elif default is not None:
                    kw.update(type=type(default))
                parser.add_argument(f"--{f.name}", default=default, **kw)
        ns = vars(parser.parse_args(args))
        return tuple(cls(**{f.name: ns[f.name] for f in dataclasses.fields(cls)}) for cls in classes)
