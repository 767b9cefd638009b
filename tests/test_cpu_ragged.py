"""BatchDetails (CORE/include/engine/batch_details.hpp:10-90): the host-side descriptor of a ragged step."""
import pytest
import torch

from proxy_inference_engine_b200.ragged import BatchDetails


def _three():
    return BatchDetails.build([(7, range(70), 0, [3, 9]),          # prefill from scratch: 70 tokens, 2 pages
                               (11, [5], 100, [1, 4]),               # decode row at position 100
                               (12, range(20), 64, [2, 8, 6])],      # prefill chunk behind 64 cached tokens
                              device="cpu")


def test_build_fills_the_reference_fields():
    b = _three()
    assert b.total_tokens_in_step == 91 == b.token_ids.numel() == b.positions.numel()
    assert b.sequence_ids == [7, 11, 12] and b.input_lengths == [70, 1, 20] and b.context_lengths == [0, 100, 64]
    assert (b.num_prefill_sequences, b.num_decode_sequences) == (2, 1)
    assert b.positions[:70].tolist() == list(range(70)) and int(b.positions[70]) == 100
    assert b.positions[71:].tolist() == list(range(64, 84))
    assert b.consolidated_block_table.shape == (3, 3) and b.consolidated_block_table.dtype == torch.int32
    assert b.consolidated_block_table[1].tolist() == [1, 4, 0]
    b.validate()
    assert b.last_token_rows().tolist() == [69, 70, 90]
    t = b.token_block_table()
    assert t.shape == (91, 3) and t[0].tolist() == [3, 9, 0] and t[70].tolist() == [1, 4, 0] and t[90].tolist() == [2, 8, 6]


@pytest.mark.parametrize("seqs", [
    [],                                            # empty batch
    [(1, [], 0, [0])],                             # a sequence without tokens
    [(1, [3], -1, [0])],                           # negative context
    [(1, [3], 64, [0])],                           # position 64 needs a second page
    [(1, [3], 0, [0]), (1, [4], 0, [1])],          # the same sequence twice
])
def test_build_rejects(seqs):
    with pytest.raises(ValueError):
        BatchDetails.build(seqs, device="cpu")


def test_validate_catches_inconsistent_descriptors():
    b = _three()
    b.total_tokens_in_step = 90
    with pytest.raises(ValueError):
        b.validate()
    b = _three()
    b.num_decode_sequences = 2
    with pytest.raises(ValueError):
        b.validate()
    b = _three()
    b.consolidated_block_table = b.consolidated_block_table[:, :1]
    with pytest.raises(ValueError):
        b.validate()
    b = _three()
    b.input_lengths = [70, 1]
    with pytest.raises(ValueError):
        b.validate()
