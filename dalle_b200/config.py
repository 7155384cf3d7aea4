"""Model hyper-parameters.

The reference hard-codes its architecture inside ``TrainingTask.__init__`` (reference task.py:62-83) and again in
``inference/run_inference.py:46-76``; here it is one dataclass with two presets:

* ``ttt``  - the production config (dim 1024, depth 64, 4+1 shared blocks, 256 text + 32x32 image tokens)
* ``tiny`` - BASELINE config 1 (dim 256, depth 2, 64 text + 16x16 image tokens) for CPU plumbing tests
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from itertools import cycle, islice
from typing import Tuple, Union

AttnId = Union[int, str]


@dataclass(frozen=True)
class DalleConfig:
    num_text_tokens: int = 32100      # tokenizer vocab (t5-small: 32100); the model adds text_seq_len pad ids
    text_seq_len: int = 256
    dim: int = 1024
    depth: int = 64
    heads: int = 16
    dim_head: int = 64
    image_fmap_size: int = 32         # VQGAN f8 on 256x256 -> 32x32 codes (reference task.py:26)
    num_image_tokens: int = 8192
    ff_mult: int = 4
    attn_types: Tuple[str, ...] = ()
    shared_attn_ids: Tuple[AttnId, ...] = ()
    shared_ff_ids: Tuple[AttnId, ...] = ()
    loss_img_weight: float = 7.0
    conv_kernel_size: int = 5
    conv_dilation: int = 1
    rotary_emb: bool = True
    reversible: bool = True
    shift_tokens: bool = True
    share_input_output_emb: bool = True
    ln_eps: float = 1e-5

    def __post_init__(self):
        depth = self.depth
        if not self.attn_types:
            types = list(islice(cycle(["axial_row", "axial_col", "axial_row", "axial_row"]), depth - 1)) + ["conv_like"]
            object.__setattr__(self, "attn_types", tuple(types))
        if not self.shared_attn_ids:
            ids = list(islice(cycle(range(4)), depth - 1)) + ["w_conv"]
            object.__setattr__(self, "shared_attn_ids", tuple(ids))
        if not self.shared_ff_ids:
            object.__setattr__(self, "shared_ff_ids", tuple(self.shared_attn_ids))
        assert len(self.attn_types) == depth and len(self.shared_attn_ids) == depth and len(self.shared_ff_ids) == depth
        seen = {}
        for t, i in zip(self.attn_types, self.shared_attn_ids):
            assert seen.setdefault(i, t) == t, f"shared attention id {i!r} used with two attention types"

    # ---- derived sizes (SURVEY §2.2 [DERIVED]) ----
    @property
    def inner_dim(self) -> int:
        return self.heads * self.dim_head

    @property
    def image_seq_len(self) -> int:
        return self.image_fmap_size ** 2

    @property
    def total_text_tokens(self) -> int:
        """text vocab incl. one unique padding id per text position."""
        return self.num_text_tokens + self.text_seq_len

    @property
    def total_tokens(self) -> int:
        return self.total_text_tokens + self.num_image_tokens

    @property
    def seq_len(self) -> int:
        return self.text_seq_len + self.image_seq_len

    @property
    def text_len(self) -> int:
        """<bos> + text tokens."""
        return self.text_seq_len + 1

    @property
    def unique_attn_ids(self) -> Tuple[AttnId, ...]:
        return tuple(dict.fromkeys(self.shared_attn_ids))

    @property
    def unique_ff_ids(self) -> Tuple[AttnId, ...]:
        return tuple(dict.fromkeys(self.shared_ff_ids))

    def layer_scale_init(self, layer_index0: int) -> float:
        d = layer_index0 + 1
        return 0.1 if d <= 18 else (1e-5 if d <= 24 else 1e-6)


def ttt_config(**overrides) -> DalleConfig:
    return replace(DalleConfig(), **overrides) if overrides else DalleConfig()


def tiny_config(**overrides) -> DalleConfig:
    base = DalleConfig(
        num_text_tokens=1000, text_seq_len=64, dim=256, depth=2, heads=4, dim_head=64, image_fmap_size=16,
        num_image_tokens=512, attn_types=("axial_row", "conv_like"), shared_attn_ids=(0, "w_conv"),
        shared_ff_ids=(0, "w_conv"))
    return replace(base, **overrides) if overrides else base


def small_config(**overrides) -> DalleConfig:
    """A few layers of every attention type with sharing - used by kernel/unit tests."""
    base = DalleConfig(
        num_text_tokens=1000, text_seq_len=64, dim=256, depth=6, heads=4, dim_head=64, image_fmap_size=16,
        num_image_tokens=512,
        attn_types=("axial_row", "axial_col", "axial_row", "axial_row", "axial_col", "conv_like"),
        shared_attn_ids=(0, 1, 2, 0, 1, "w_conv"), shared_ff_ids=(0, 1, 2, 0, 1, "w_conv"))
    return replace(base, **overrides) if overrides else base


PRESETS = {"ttt": ttt_config, "tiny": tiny_config, "small": small_config}


def get_config(name: str, **overrides) -> DalleConfig:
    return PRESETS[name](**overrides)
