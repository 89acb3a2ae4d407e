"""Numerical constants of the hot path. Same names and values as the reference's `mantrap/constants.py`
(cited line numbers) so that host code written against the reference reads the same here."""
ID_EGO = "ego"                                   # :4

PED_SPEED_MAX = 2.5                              # :28 [m/s]
PED_ACC_MAX = 20.0                               # :30 (not used by the reference either)
ROBOT_SPEED_MAX = 2.0                            # :30 [m/s]
ROBOT_ACC_MAX = 2.0                              # :31 [m/s^2]

ENV_DT_DEFAULT = 0.4                             # :39
ENV_X_AXIS_DEFAULT = (-10, 10)                   # :40
ENV_Y_AXIS_DEFAULT = (-10, 10)                   # :41
ENV_VAR_INITIAL = 1e-6                           # :43
ENV_NUM_PARTICLES = 5                            # :45
ENV_PARTICLE_NOISE = 1e-6                        # :46
KALMAN_ADDITIVE_NOISE = 0.2                      # :48
SOCIAL_FORCES_DEFAULT_TAU = 0.4                  # :50
SOCIAL_FORCES_DEFAULT_V0 = 4.0, 1.0              # :51 (mean, "variance" used as std)
SOCIAL_FORCES_DEFAULT_SIGMA = 0.9, 0.3           # :52
SOCIAL_FORCES_MAX_GOAL_DISTANCE = 0.3            # :53
SOCIAL_FORCES_MAX_INTERACTION_DISTANCE = 2.0     # :54
POTENTIAL_FIELD_V0_DEFAULT = 2.0, 1.0            # :56
POTENTIAL_FIELD_MAX_THETA = 30.0                 # :57 [deg]

SOLVER_HORIZON_DEFAULT = 5                       # :68
SOLVER_CONSTRAINT_LIMIT = 1e-3                   # :69
SOLVER_GOAL_END_DISTANCE = 0.5                   # :70
IPOPT_MAX_CPU_TIME_DEFAULT = 2.0                 # :81
IPOPT_OPTIMALITY_TOLERANCE = 0.1                 # :82
SEARCH_MAX_CPU_TIME = 0.5                        # :86
MCTS_NUMBER_BREADTH_SAMPLES = 2                  # :87
MCTS_NUMBER_DEPTH_SAMPLES = 2                    # :88

OBJECTIVE_PROB_INTERACT_MAX = 120.0              # :100
OBJECTIVE_POS_INTERACT_MAX = 3.0                 # :101
OBJECTIVE_VEL_INTERACT_MAX = 1.0                 # :102
OBJECTIVE_ACC_INTERACT_MAX = 1.0                 # :103
CONSTRAINT_MIN_L2_DISTANCE = 0.5                 # :107
CONSTRAINT_VIOLATION_PRECISION = 1e-5            # :108
CONSTRAINT_ELLIPSOID_NUM_MODES = 5               # :109
CONSTRAINT_ELLIPSOID_MIN_SCALE = 0.01            # modules/baselines/ellipsoid_interact.py:52 (literal in the reference)
ATTENTION_EUCLIDEAN_RADIUS = 4.0                 # :111
ATTENTION_CLOSEST_RADIUS = 4.0                   # :112

TAG_OPTIMIZATION = "optimization"                # :16
TAG_WARM_START = "warm_start"                    # :17
WARM_START_HARD = "hard"                         # :72
WARM_START_ENCODING = "encoding"                 # :72
WARM_START_SOFT = "soft"                         # :73 (needs HJReachabilityModule: not on this path)
WARM_START_POTENTIAL = "potential"               # :74
WARM_START_ZEROS = "zeros"                       # :76
WARM_START_PRE_COMPUTATION_NUM = 100             # :77
WARM_START_PRE_COMPUTATION_HORIZON = 10          # :78
WARM_START_PRE_COMPUTATION_FILE = ("encoding.pt", "solution.pt")   # :79 (third_party/warm_start/ of the reference)
