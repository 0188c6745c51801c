"""Drop-in for the reference's config.py (config.py:1-27): same names, same values, consumed
through ``from config import *`` by core / cbf / controller / train.  The values that reach the
CUDA kernels are passed per call through the C ABI (include/macbf_b200.h); nothing is baked in
at compile time.
"""

# --- simulation / CBF condition (config.py:1-4) ---
TIME_STEP = 0.1            # Euler step and derivative condition (core.py:134,139; train.py:70)
ALPHA_CBF = 1.0            # class-K gain in h' - h + dt*alpha*h (core.py:139)

# --- geometry (config.py:6-11) ---
DIST_MIN_THRES = 0.07      # safety radius r (core.py:197, cbf.py:29, train.py:51)
DIST_MIN_CHECK = 0.07      # evaluation-time radius and rollout early-stop (core.py:219, train.py:80)
DIST_MIN_ENLARGED = 0.07   # unused by the reference
OBS_RADIUS = 1.0           # sensing radius of the masks (cbf.py:36, controller.py:39)
TOP_K = 12                 # neighbours per agent, self included (core.py:191, cbf.py:31, controller.py:35)

# --- time to collision (config.py:13-14) ---
TIME_TO_COLLISION = 2.0         # training mask horizon (core.py:204-205)
TIME_TO_COLLISION_CHECK = 0.1   # evaluation mask horizon (core.py:226-227)

# --- optimisation (config.py:2,16-24) ---
LEARNING_RATE = 1e-4
WEIGHT_DECAY = 1e-8
TRAIN_STEPS = 70000
INNER_LOOPS = 50
EVALUATE_STEPS = 10        # unused by the reference
REFINE_LOOPS = 50          # unused by the reference
REFINE_LEARNING_RATE = 0.3  # unused by the reference
DISPLAY_STEPS = 200        # unused by the reference
SAVE_STEPS = 1000          # unused by the reference

# --- action noise during rollout (config.py:26-27; probability 0 => dead branch) ---
ADD_NOISE_PROB = 0.0
NOISE_SCALE = 0.3
