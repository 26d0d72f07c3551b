"""Headless stand-in for the reference's `game` module (src/game/mod.rs): the CALLER of the render
boundary.  Same accessors (borrow_camera, borrow_world_mut, get_sun_angle, tick) so that
`Pipeline.draw_frame(game)` is driven exactly as src/bin/main.rs:41-54 drives it."""
import numpy as np

from .render import Camera, compute_triple_euler_vector
from .world import ChunkStorage


class ControlSet:
    """src/game/control.rs: named held/pressed state; keys are plain strings here (no winit)."""
    NAMES = ("up", "down", "left", "right", "forward", "backward", "sunup", "sundown")

    def __init__(self):
        self.this_state = {n: False for n in self.NAMES}
        self.last_state = dict(self.this_state)

    def tick(self):
        self.last_state = dict(self.this_state)

    def on_pressed(self, name):
        if name in self.this_state:
            self.this_state[name] = True

    def on_released(self, name):
        if name in self.this_state:
            self.this_state[name] = False

    def is_held(self, name):
        return self.this_state.get(name, False)

    def is_pressed(self, name):
        return self.this_state.get(name, False) and not self.last_state.get(name, False)

    def is_released(self, name):
        return not self.this_state.get(name, False) and self.last_state.get(name, False)


class Game:
    def __init__(self, args=None, world=None):
        """mod.rs:37-58: six optional positional values (x y z heading pitch sun_angle)."""
        self.camera = Camera()
        self.world = world if world is not None else ChunkStorage()
        self.controls = ControlSet()
        self.sun_angle = 0.0
        if args:
            self.camera.origin = [float(args[0]), float(args[1]), float(args[2])]
            self.camera.heading, self.camera.pitch, self.sun_angle = float(args[3]), float(args[4]), float(args[5])
        else:
            self.camera.origin = [-30.0, -128.0, 100.0]

    def tick(self, dt):
        """mod.rs:61-96."""
        c = self.controls
        if c.is_held("sunup"):
            self.sun_angle += dt
        elif c.is_held("sundown"):
            self.sun_angle -= dt
        dx = -1.0 if c.is_held("left") else 1.0 if c.is_held("right") else 0.0
        dy = -1.0 if c.is_held("backward") else 1.0 if c.is_held("forward") else 0.0
        dz = -1.0 if c.is_held("down") else 1.0 if c.is_held("up") else 0.0
        amount = np.float32(dt * 50.0)
        fwd, up, right = compute_triple_euler_vector(self.camera.heading, self.camera.pitch)
        fwd, up, right = (v / np.sqrt((v * v).sum(dtype=np.float32)) for v in (fwd, up, right))
        o = np.asarray(self.camera.origin, np.float32) + amount * fwd * np.float32(dy) + amount * up * np.float32(dz) \
            + amount * right * np.float32(dx)
        self.camera.origin = [float(v) for v in o]

    def borrow_world(self):
        return self.world

    def borrow_world_mut(self):
        return self.world

    def borrow_camera(self):
        return self.camera

    def borrow_controls(self):
        return self.controls

    def borrow_controls_mut(self):
        return self.controls

    def get_sun_angle(self):
        return self.sun_angle
