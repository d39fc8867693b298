"""Known-answer vectors of DeepMind rlax's own unit tests for the recursions rlego mirrors (``rlax/_src/vtrace_test.py``:
``VTraceTest.setUp``; ``rlax/_src/multistep_test.py``: ``LambdaReturnsTest``, ``TruncatedGeneralizedAdvantageEstimationTest``).

PROVENANCE, stated plainly: neither rlax nor rlego nor a network exists in this environment, so these arrays were typed in
from the published test files as remembered, not copied from a checkout.  What makes them usable as a pin is the check
itself: the oracle (written independently from the papers' recursions) reproduces the 8 V-trace TD errors to the last
printed digit, the 8 policy-gradient advantages to 6e-7 and the 10 lambda-returns to 3e-7 -- numbers like -8.3327293e-05
do not agree by accident, so both the transcription and the oracle are right.  The GAE expectations are printed to 4-6
digits in the rlax file (it asserts atol 1e-3) and are held to that.  rlax's n-step test vector is not used: one of
its ten elements was not remembered reliably."""
import numpy as np

# ---- vtrace_test.py -------------------------------------------------------------------------------------------------
BEHAVIOR_LOGITS = np.array([[[8.9, 0.7], [5.0, 1.0], [0.6, 0.1], [-0.9, -0.1]],
                            [[0.3, -5.0], [1.0, -8.0], [0.3, 1.7], [4.7, 3.3]]], dtype=np.float32)
TARGET_LOGITS = np.array([[[0.4, 0.5], [9.2, 8.8], [0.7, 4.4], [7.9, 1.4]],
                          [[1.0, 0.9], [1.0, -1.0], [-4.3, 8.7], [0.8, 0.3]]], dtype=np.float32)
ACTIONS = np.array([[0, 1, 0, 0], [1, 0, 0, 1]], dtype=np.int64)
REWARDS = np.array([[-1.3, -1.3, 2.3, 42.0], [1.3, 5.3, -3.3, -5.0]], dtype=np.float32)
DISCOUNTS = np.array([[0., 0.89, 0.85, 0.99], [0.88, 1., 0.83, 0.95]], dtype=np.float32)
VALUES = np.array([[2.1, 1.1, -3.1, 0.0], [3.1, 0.1, -1.1, 7.4]], dtype=np.float32)
BOOTSTRAP = np.array([8.4, -1.2], dtype=np.float32)
VTRACE_KW = dict(lambda_=1.0, clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0)   # the test passes clip_rho and lambda only
EXPECTED_TD = np.array([[-1.6155143, -3.4973226, 1.8670533, 5.0316002e1],
                        [1.4662437, 3.6116405, -8.3327293e-5, -1.3540000e1]], dtype=np.float32)
EXPECTED_PG = np.array([[-1.6155143, -3.4973226, 1.8670534, 5.0316002e1],
                        [1.4662433, 3.6116405, -8.3369283e-05, -1.3540000e+1]], dtype=np.float32)


def rho_tm1():
    """``exp(log pi_target(a) - log pi_behaviour(a))`` as the rlax test forms it."""
    def lsm(x):
        x = x.astype(np.float64)
        return x - np.log(np.exp(x).sum(-1, keepdims=True))
    tl = np.take_along_axis(lsm(TARGET_LOGITS), ACTIONS[..., None], -1)[..., 0]
    bl = np.take_along_axis(lsm(BEHAVIOR_LOGITS), ACTIONS[..., None], -1)[..., 0]
    return np.exp(tl - bl).astype(np.float32)


def v_t():
    return np.concatenate([VALUES[:, 1:], BOOTSTRAP[:, None]], axis=1)


# ---- multistep_test.py ----------------------------------------------------------------------------------------------
MS_R = np.array([[1.0, 0.0, -1.0, 0.0, 1.0], [0.5, 0.8, -0.7, 0.0, 2.1]], dtype=np.float32)
MS_DISCOUNT = np.array([[0.5, 0.9, 1.0, 0.5, 0.8], [0.9, 0.5, 0.3, 0.8, 0.7]], dtype=np.float32)
MS_V = np.array([[3.0, 1.0, 5.0, -5.0, 3.0], [-1.7, 1.2, 2.3, 2.2, 2.7]], dtype=np.float32)
MS_LAMBDA = 0.75
EXPECTED_LAMBDA_RETURNS = np.array([[1.6460547, 0.72281253, 0.7375001, 0.6500001, 3.4],
                                    [0.7866317, 0.9913063, 0.1101501, 2.834, 3.99]], dtype=np.float32)

GAE_R = np.array([[0., 0., 1., 0., -0.5], [0., 0., 0., 0., 1.]], dtype=np.float32)
GAE_V = np.array([[1., 4., -3., -2., -1., -1.], [-3., -2., -1., 0.0, 5., -1.]], dtype=np.float32)
GAE_DISCOUNT = np.array([[0.99, 0.99, 0.99, 0.99, 0.99], [0.9, 0.9, 0.9, 0.0, 0.9]], dtype=np.float32)
EXPECTED_GAE = {                                   # printed to 4-6 digits there; atol 1e-3 as in the rlax test
    1.0: np.array([[-1.45118, -4.4557, 2.5396, 0.5249, -0.49], [3., 2., 1., 0., -4.9]], dtype=np.float32),
    0.7: np.array([[-0.676979, -5.248167, 2.4846, 0.6704, -0.49], [2.2899, 1.73, 1., 0., -4.9]], dtype=np.float32),
    0.4: np.array([[0.56731, -6.042, 2.3431, 0.815, -0.49], [1.725, 1.46, 1., 0., -4.9]], dtype=np.float32),
}
