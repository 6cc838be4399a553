class _L:
    def __init__(self, *a, **k):
        pass


TensorBoardLogger = CSVLogger = NeptuneLogger = _L
