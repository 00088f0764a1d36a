"""Error type compatible with seldon_core's SeldonMicroserviceException (python/seldon_core/flask_utils.py:67-81):
same attributes (.message, .status_code, .reason) and .to_dict() shape, so the wrapper's error handler renders it."""


class SeldonB200Error(Exception):
    status_code = 400

    def __init__(self, message, status_code=None, payload=None, reason="MICROSERVICE_BAD_DATA"):
        Exception.__init__(self, message)
        self.message = message
        if status_code is not None:
            self.status_code = status_code
        self.payload = payload
        self.reason = reason

    def to_dict(self):
        return {"status": {"status": 1, "info": self.message, "code": -1, "reason": self.reason}}
