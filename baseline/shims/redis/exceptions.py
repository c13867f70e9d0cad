class RedisError(Exception):
    pass
