"""Procedural table image.

The reference's table is a photographic sprite that lived in the missing `air_hockey` submodule
(the only surviving picture is /root/reference/images/simulation.png: rounded light
playfield on a dark surround, red centre line with a gap and a dot, four red circles near the corners, dotted
air-hole columns).  This draws a stand-in with the same elements.  It is DATA handed to the engine through
`ahk_set_table_image` (any uint8[700][450][3] works); integer arithmetic only, so it is identical everywhere.
"""
import numpy as np

W, H = 450, 700
RINK_L, RINK_R, RINK_T, RINK_B = 25, 425, 25, 675
CORNER_R = 60
GOAL_L, GOAL_R = 145, 305
MID_Y = 350


def _hash_noise(x, y):
    """Deterministic per-pixel texture in [0, 15] (32-bit integer mix), so that the image has no flat areas."""
    h = (x.astype(np.uint64) * np.uint64(0x9E3779B1) + y.astype(np.uint64) * np.uint64(0x85EBCA77)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    return (h & np.uint64(15)).astype(np.int32)


def make_table_image():
    """-> uint8[700][450][3], RGB."""
    y, x = np.mgrid[0:H, 0:W].astype(np.int32)
    noise = _hash_noise(x, y)

    img = np.empty((H, W, 3), dtype=np.int32)
    # dark surround
    img[..., 0] = 28 + (noise >> 1)
    img[..., 1] = 30 + (noise >> 1)
    img[..., 2] = 36 + (noise >> 1)

    # rounded playfield
    inside = (x >= RINK_L) & (x <= RINK_R) & (y >= RINK_T) & (y <= RINK_B)
    for cx, cy, sx, sy in ((RINK_L + CORNER_R, RINK_T + CORNER_R, -1, -1), (RINK_R - CORNER_R, RINK_T + CORNER_R, 1, -1),
                           (RINK_L + CORNER_R, RINK_B - CORNER_R, -1, 1), (RINK_R - CORNER_R, RINK_B - CORNER_R, 1, 1)):
        in_square = ((x - cx) * sx > 0) & ((y - cy) * sy > 0)
        inside &= ~(in_square & ((x - cx) ** 2 + (y - cy) ** 2 > CORNER_R ** 2))
    shade = (y * 12) // H  # slight vertical gradient, like the photo
    field = np.stack([196 + noise - shade, 198 + noise - shade, 190 + noise - shade], axis=-1)
    img[inside] = field[inside]

    red = np.array([170, 40, 50], dtype=np.int32)

    # goal mouths (dark slots in the top and bottom walls)
    mouth = (x >= GOAL_L) & (x <= GOAL_R) & (((y >= RINK_T - 12) & (y < RINK_T)) | ((y > RINK_B) & (y <= RINK_B + 12)))
    img[mouth] = np.array([8, 8, 10], dtype=np.int32)

    # centre line with a gap and a dot
    line = inside & (np.abs(y - MID_Y) <= 2) & (np.abs(x - W // 2) >= 40)
    img[line] = red
    dot = (x - W // 2) ** 2 + (y - MID_Y) ** 2 <= 36
    img[dot] = red

    # four face-off circles
    for cx, cy in ((115, 130), (335, 130), (115, 570), (335, 570)):
        d2 = (x - cx) ** 2 + (y - cy) ** 2
        ring = (d2 <= 38 ** 2) & (d2 >= 34 ** 2)
        img[ring] = red

    # air holes: dotted columns
    holes = inside & (x % 25 == 12) & (y % 14 == 7) & (np.abs(y - MID_Y) > 6)
    img[holes] = np.array([120, 122, 118], dtype=np.int32)

    return np.clip(img, 0, 255).astype(np.uint8)
