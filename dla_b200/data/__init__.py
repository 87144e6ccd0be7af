"""Data layer (parity: ``R/src/data/datasets.py``)."""
from dla_b200.data.datasets import (  # noqa: F401
    EvalPromptDataset, InstructionDataset, PackedCollator, PreferenceDataset, Sample,
    TeacherRolloutDataset, build_instruction_dataset, build_preference_dataset,
    load_instruction_records, load_preference_records, pad_batch, read_jsonl, write_jsonl,
)
from dla_b200.data.sampler import RankSampler, get_distributed_sampler, split_between_processes  # noqa: F401
from dla_b200.data.tokenizer import ByteTokenizer, load_tokenizer  # noqa: F401
