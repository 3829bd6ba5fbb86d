"""Drop-in for the reference's ``robot`` module (/root/reference/robot.py:3-14): the pose holder
``System`` reads its state dimension and initial state from."""
import numpy as np


class Robot:
    def __init__(self):
        self.pos = np.zeros(3)            # (x, y, theta), mm / rad
        self.dimX = int(self.pos.shape[0])
        self.name = "Main robot"

    def get_dim_x(self):
        return self.dimX

    def get_pos(self):
        return self.pos
