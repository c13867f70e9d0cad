class PyMongoError(Exception):
    pass
