"""User-level scenarios on the CPU (oracle) path; the same bodies run on the CUDA path in test_gpu_scenarios.py."""
import scenarios


def test_resnet_train_eval_and_checkpoint_restore(tmp_path):
    scenarios.resnet_train_eval_and_cross_device_restore(tmp_path, gpu=False)


def test_placeholder_and_datapath_loaders_learn(tmp_path):
    scenarios.loader_kinds_learn_the_toy_problem(tmp_path, gpu=False)


def test_adam_follows_a_decaying_learning_rate_handler(tmp_path):
    scenarios.adam_with_decaying_learning_rate(tmp_path, gpu=False)


def test_two_in_process_towers_equal_one_tower_on_the_concatenated_batch():
    scenarios.in_process_towers_equal_one_big_batch("cpu")
