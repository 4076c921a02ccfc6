"""Board geometry, camera and image constants (values of reference ``language_table/environments/constants.py:22-65``)."""
import math

import numpy as np

EFFECTOR_HEIGHT = 0.145
X_MIN, X_MAX = 0.15, 0.6
Y_MIN, Y_MAX = -0.3048, 0.3048
CENTER_X = (X_MAX - X_MIN) / 2.0 + X_MIN
CENTER_Y = (Y_MAX - Y_MIN) / 2.0 + Y_MIN
WORKSPACE_BOUNDS_BUFFER = 0.08
BLOCK_DISTANCE_THRESHOLD = 0.0175
ARM_DISTANCE_THRESHOLD = 0.06
INSTRUCTION_LENGTH = 512            # bytes of the zero-padded utf-8 instruction in the observation
WORKSPACE_BOUNDS = np.array(((X_MIN, Y_MIN), (X_MAX, Y_MAX)))

CAMERA_POSE = (0.75, 0.0, 0.5)
CAMERA_ORIENTATION = (math.pi / 5, math.pi, -math.pi / 2)      # intrinsic x-y-z Euler angles, as PyBullet's getQuaternionFromEuler
IMAGE_WIDTH, IMAGE_HEIGHT = 320, 180
CAMERA_INTRINSICS = (0.803 * IMAGE_WIDTH, 0, IMAGE_WIDTH / 2.0, 0, 0.803 * IMAGE_WIDTH, IMAGE_HEIGHT / 2.0, 0, 0, 1)

# kinematic stand-in for the Bullet contact model (not in the reference: it simulates rigid bodies)
EFFECTOR_RADIUS = 0.0125
BLOCK_RADIUS = 0.02
BLOCK_HEIGHT = 0.02
CONTROL_HZ = 10.0
