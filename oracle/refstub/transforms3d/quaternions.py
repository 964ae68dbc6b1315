rotate_vector = qconjugate = mat2quat = qmult = None
