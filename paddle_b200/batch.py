"""paddle.batch (reference: python/paddle/batch.py): wrap a sample reader into a reader of lists of samples."""

__all__ = []


def batch(reader, batch_size, drop_last=False):
    batch_size = int(batch_size)
    if batch_size <= 0:
        raise ValueError(f"batch_size should be a positive integer value, but got batch_size={batch_size}")

    def batch_reader():
        chunk = []
        for instance in reader():
            chunk.append(instance)
            if len(chunk) == batch_size:
                yield chunk
                chunk = []
        if chunk and not drop_last:
            yield chunk
    return batch_reader
