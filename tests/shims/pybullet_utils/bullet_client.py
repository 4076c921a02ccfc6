"""Test-harness stand-in: see ../README.md."""


class BulletClient:
    def __init__(self, connection_mode=None, **kwargs):
        self.connection_mode = connection_mode

    def resetSimulation(self, *a, **k):
        pass

    def configureDebugVisualizer(self, *a, **k):
        pass

    def setPhysicsEngineParameter(self, *a, **k):
        pass

    def disconnect(self, *a, **k):
        pass
