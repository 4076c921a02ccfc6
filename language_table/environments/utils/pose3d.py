"""Minimal rigid transform (reference ``environments/utils/pose3d.py:41``): a rotation (scipy ``Rotation``) + translation."""
import dataclasses

import numpy as np
from scipy.spatial import transform


class NoCopyAsDict:
    """Mixin for dataclasses: ``asdict()`` without the deep copy ``dataclasses.asdict`` performs (reference ``pose3d.py:24-37``)."""

    def asdict(self):
        return {f.name: getattr(self, f.name) for f in dataclasses.fields(self)}


@dataclasses.dataclass
class Pose3d(NoCopyAsDict):
    rotation: transform.Rotation
    translation: np.ndarray

    @property
    def vec7(self):
        return np.concatenate([self.translation, self.rotation.as_quat()])

    def serialize(self):
        return {"rotation": self.rotation.as_quat().tolist(), "translation": np.asarray(self.translation).tolist()}

    @staticmethod
    def deserialize(data):
        return Pose3d(rotation=transform.Rotation.from_quat(data["rotation"]), translation=np.array(data["translation"]))

    def __eq__(self, other):
        return np.allclose(self.rotation.as_quat(), other.rotation.as_quat()) and np.allclose(self.translation, other.translation)
