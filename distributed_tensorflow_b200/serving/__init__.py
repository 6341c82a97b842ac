from .server import ModelServer, make_http_server, main  # noqa: F401
