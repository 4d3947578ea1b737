"""Net wrapper: forward "aspect" around a Model. Parity: /root/reference/distribute_net.py:9-38.

`process(x) = post_process(inference(pre_process(x)))`; users override `pre_process` / `post_process`.
"""


def current_model(**kwargs):
    """Local decorator of the reference (key `net`), distribute_net.py:9-15."""

    def deco(obj):
        if "net" in kwargs:
            setattr(obj, "net", kwargs["net"])
        return obj

    return deco


class Net(object):
    def __init__(self, model):
        self.model = model

    def inference(self, pre_processed_data, *args, **kwargs):
        return self.model.inference(pre_processed_data)

    def pre_process(self, input_data, *args, **kwargs):
        return input_data

    def post_process(self, inference_result, *args, **kwargs):
        return inference_result

    def process(self, input_data, *args, **kwargs):
        pre = self.pre_process(input_data, *args, **kwargs)
        out = self.inference(pre)
        return self.post_process(out, *args, **kwargs)
