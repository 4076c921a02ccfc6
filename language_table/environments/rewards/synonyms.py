"""Referring expressions and verb/preposition vocabularies of the structured board language
(reference ``rewards/synonyms.py``)."""
import collections


def get_color_shape(block):
    color, shape = block.split("_")
    return color, shape


def count_color_shape(blocks_on_table):
    colors, shapes = zip(*(get_color_shape(b) for b in blocks_on_table))
    return collections.Counter(colors), collections.Counter(shapes)


def get_block_synonyms(block, blocks_on_table):
    """Unambiguous ways to name ``block`` given what else is on the table: "<colour> block" if the colour is unique,
    "<shape>" if the shape is unique, and always "<colour> <shape>"."""
    color, shape = get_color_shape(block)
    n_color, n_shape = count_color_shape(blocks_on_table)
    names = []
    if n_color[color] == 1:
        names.append(f"{color} block")
    if n_shape[shape] == 1:
        names.append(shape)
    names.append(f"{color} {shape}")
    return names


PUSH_VERBS = ["push the", "move the", "slide the", "put the"]
PREPOSITIONS = ["to the", "towards the", "close to the", "next to the"]
POINT_PREPOSITIONS = [
    "point next to the", "point close to the", "point to the", "point at the", "move the arm next to the",
    "move the arm close to the", "move the arm to the", "move your arm next to the", "move your arm close to the",
    "move your arm to the", "move next to the", "move close to the", "move to the",
]
