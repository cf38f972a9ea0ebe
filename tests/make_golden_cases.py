"""Problems whose oracle outputs are committed in tests/golden/oracle_golden.npz."""
import cases

INPUT_SEED = 20261104


def problems():
    return {
        # BASELINE.json configs[0]: B=1 H=4 S=128 D=64 non-causal
        "c1_b1h4s128d64": cases.dense(1, 128, 128, 4, 4, 64),
        "causal_gqa_drop": cases.dense(2, 96, 96, 4, 2, 64, is_causal=True, p_dropout=0.1, seed=1234, offset=0),
        "d128_fp16": cases.dense(1, 80, 112, 2, 2, 128, is_causal=True, is_bf16=False),
        "varlen_mixed": cases.varlen([5, 64, 1, 33], [7, 64, 9, 33], 4, 2, 64, is_causal=True,
                                     p_dropout=0.2, seed=3, offset=2 ** 33 + 1),
    }
