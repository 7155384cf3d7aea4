"""``TrainingTask``: builds model, tokenizer, DHT, collaborative optimizer and data lazily (reference task.py:45-181).

Same public attributes as the reference (``model``, ``tokenizer``, ``authorizer``, ``dht``,
``collaborative_optimizer``, ``training_dataset``, ``data_collator``, ``validators``, ``local_public_key``); the
model is this repository's ``DALLE`` with the TTT configuration, the optimizer this repository's LAMB-8bit behind
this repository's collaborative ``Optimizer``.
"""
from __future__ import annotations

import hashlib
import logging
import os
import socket
from dataclasses import asdict
from pathlib import Path

import torch

import utils
from arguments import BasePeerArguments, CollaborativeArguments, HFTrainerArguments
from data import PadCollator, make_dataset
from dalle_b200.collab import Float16Compression, Optimizer, SizeAdaptiveCompression, Uniform8BitQuantization
from dalle_b200.config import get_config
from dalle_b200.dht import DHT
from dalle_b200.models.dalle import DALLE, ModelWrapper
from dalle_b200.optim.lamb8bit import LAMB8Bit, get_linear_schedule_with_warmup
from dalle_b200.trainer import param_groups
from dalle_b200.training.tokenizer import load_tokenizer

logger = logging.getLogger(__name__)


class VQGanParams:
    """Parameter-free stand-in for the VQGAN: only the geometry the model needs (reference task.py:25-32)."""

    def __init__(self, *, num_layers=3, image_size=256, num_tokens=8192, is_gumbel=True):
        self.num_layers, self.image_size, self.num_tokens, self.is_gumbel = num_layers, image_size, num_tokens, is_gumbel


def group_rank_world():
    """(rank, world) of the multi-GPU peer this process belongs to: a torchrun launch is ONE swarm peer made of one
    process per GPU; a plain ``python run_trainer.py`` is a peer of its own, (0, 1)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class TrainingTask:
    _authorizer = _dht = _collaborative_optimizer = _training_dataset = None

    def __init__(self, peer_args: BasePeerArguments, trainer_args: HFTrainerArguments,
                 collab_args: CollaborativeArguments):
        self.peer_args, self.trainer_args, self.collab_args = peer_args, trainer_args, collab_args
        self.group_rank, self.group_world = group_rank_world()
        # the reference dereferences authorizer.username unconditionally and crashes with --authorize false
        auth = self.authorizer
        self.trainer_args.run_name = auth.username if auth is not None else (
            trainer_args.run_name or f"{socket.gethostname()}-{os.getpid()}")
        self.validators, self.local_public_key = utils.make_validators(peer_args.experiment_prefix)
        torch.manual_seed(trainer_args.seed)
        self.tokenizer = load_tokenizer(peer_args.tokenizer_path)

        vae = VQGanParams()
        preset = getattr(trainer_args, "model_preset", "ttt")
        if preset == "ttt":
            cfg = get_config("ttt", num_text_tokens=self.tokenizer.vocab_size, text_seq_len=trainer_args.text_seq_length,
                             image_fmap_size=vae.image_size // 2 ** vae.num_layers, num_image_tokens=vae.num_tokens)
        else:
            cfg = get_config(preset)
            trainer_args.text_seq_length = cfg.text_seq_len
        logger.info("Creating model")
        dalle = DALLE(cfg, seed=trainer_args.seed)
        dalle.vae = vae
        logger.info(f"Trainable parameters: {sum(p.numel() for p in dalle.parameters() if p.requires_grad)}")
        self.model = ModelWrapper(dalle)

        out_dir = Path(trainer_args.output_dir)
        ckpts = sorted(out_dir.glob("checkpoint*"), key=os.path.getctime)
        logger.info(f"Checkpoint dir {out_dir}, contents {ckpts}")
        if ckpts:
            logger.info(f"Loading model from {ckpts[-1]}")
            self.model.load_state_dict(torch.load(ckpts[-1] / "model_state.pt", map_location="cpu"))

    # ---- lazily built collaborators ----------------------------------------------------------------
    @property
    def authorizer(self):
        if self._authorizer is None and self.peer_args.authorize:
            from huggingface_auth import authorize_with_huggingface

            self._authorizer = authorize_with_huggingface()
        return self._authorizer

    @property
    def dht(self) -> DHT:
        if self.group_rank != 0:
            return None   # inside a multi-GPU peer only the leader rank talks to the swarm
        if self._dht is None:
            a = self.peer_args
            self._dht = DHT(start=True, initial_peers=a.initial_peers, client_mode=a.client_mode,
                            host_maddrs=a.host_maddrs, announce_maddrs=a.announce_maddrs, use_ipfs=a.use_ipfs,
                            record_validators=self.validators, identity_path=a.identity_path,
                            authorizer=self.authorizer)
            if a.client_mode:
                logger.info(f"Created client mode peer with peer_id={self._dht.peer_id}")
            else:
                utils.log_visible_maddrs(self._dht.get_visible_maddrs(), only_p2p=a.use_ipfs)
        return self._dht

    @property
    def collaborative_optimizer(self) -> Optimizer:
        if self._collaborative_optimizer is None:
            params, make_opt, make_sched = self._get_local_optimizer_and_scheduler(self.trainer_args)
            compression = SizeAdaptiveCompression(threshold=2 ** 16 + 1, less=Float16Compression(),
                                                  greater_equal=Uniform8BitQuantization())
            if self.group_world > 1:
                # one swarm peer = all ranks of this torchrun launch: NVLink-fused averaging + sharded LAMB inside the
                # group, the leader rank on the swarm's TCP plane (dalle_b200/collab/group_optimizer.py)
                from dalle_b200.collab.group_optimizer import GroupOptimizer

                a = self.trainer_args
                per_rank = a.per_device_train_batch_size * a.gradient_accumulation_steps
                self._collaborative_optimizer = GroupOptimizer(
                    dht=self.dht, run_id=self.peer_args.experiment_prefix, params=params, optimizer=make_opt,
                    scheduler=make_sched, delay_optimizer_step=True, batch_size_per_step=per_rank * self.group_world,
                    grad_compression=compression, state_averaging_compression=compression,
                    client_mode=self.peer_args.client_mode, verbose=True,
                    target_group_size=int(os.environ.get("DALLE_B200_TARGET_GROUP_SIZE", "0")) or None,
                    **asdict(self.collab_args))
                return self._collaborative_optimizer
            self._collaborative_optimizer = Optimizer(
                dht=self.dht, run_id=self.peer_args.experiment_prefix, params=params, optimizer=make_opt,
                scheduler=make_sched, offload_optimizer=True, delay_grad_averaging=False, delay_optimizer_step=True,
                batch_size_per_step=self.trainer_args.batch_size_per_step, grad_compression=compression,
                state_averaging_compression=compression, client_mode=self.peer_args.client_mode, verbose=True,
                auxiliary=bool(getattr(self, "auxiliary", False)),
                **asdict(self.collab_args))
        return self._collaborative_optimizer

    def _get_local_optimizer_and_scheduler(self, a: HFTrainerArguments):
        params = param_groups(self.model.model, a.weight_decay)
        make_opt = lambda p: LAMB8Bit(p, lr=a.learning_rate, betas=(a.adam_beta1, a.adam_beta2), eps=a.adam_epsilon,
                                      weight_decay=a.weight_decay, max_grad_norm=a.max_grad_norm,
                                      clamp_value=a.clamp_value, reuse_grad_buffers=True)
        make_sched = lambda opt: get_linear_schedule_with_warmup(opt, a.warmup_steps, a.total_steps)
        return params, make_opt, make_sched

    @property
    def training_dataset(self):
        if self._training_dataset is None:
            # deterministic per-peer shuffle seed (the reference's hash(bytes) is salted per process)
            # (inside a multi-GPU peer every rank has its own key, hence its own order)
            seed = int.from_bytes(hashlib.sha256(self.local_public_key).digest()[:4], "little") % 2 ** 31
            cfg = self.model.model.cfg
            self._training_dataset = make_dataset(
                self.tokenizer, shuffle_seed=seed, max_sequence_length=self.trainer_args.text_seq_length,
                synthetic=getattr(self.trainer_args, "synthetic_data", False), image_tokens=cfg.image_seq_len,
                image_vocab=cfg.num_image_tokens)
        return self._training_dataset

    @property
    def data_collator(self):
        return PadCollator(self.tokenizer.pad_token_id, self.trainer_args.text_seq_length)
