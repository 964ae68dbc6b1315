normalized_vector = None
