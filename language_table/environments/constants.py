"""Board geometry, camera and image constants (values of reference ``language_table/environments/constants.py:22-65``)."""
import math

import numpy as np

from scipy.spatial import transform

EFFECTOR_HEIGHT = 0.145
EFFECTOR_DOWN_ROTATION = transform.Rotation.from_rotvec([0, math.pi, 0])
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

# xArm6 joint angles whose forward kinematics is rotation rotvec [0, pi, 0], translation [0.3, -0.2, 0.145]
INITIAL_JOINT_POSITIONS = np.array([-0.5875016909413221, 0.15985553866983415, -0.4992862770497537, 0.0017427885915130214,
                                    0.33927183830553914, -3.7249551487437524])
STEP_FREQUENCY = 240.0             # simulation sub-steps per second (Bullet's default time step)

# kinematic stand-in for the Bullet contact model (not in the reference: it simulates rigid bodies)
EFFECTOR_RADIUS = 0.0125
BLOCK_RADIUS = 0.02
BLOCK_HEIGHT = 0.02
CONTROL_HZ = 10.0

# asset paths in the reference's convention (``constants.py:22,42``): Bullet's ground plane and the board
PLANE_URDF_PATH = "third_party/bullet/examples/pybullet/gym/pybullet_data/plane.urdf"
WORKSPACE_URDF_PATH = "third_party/py/language_table/environments/assets/workspace_real.urdf"
