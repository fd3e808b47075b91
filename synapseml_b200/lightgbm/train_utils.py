"""TrainUtils: mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/TrainUtils.scala
(createBooster :16-29, updateOneIteration :77-96, executeTrainingIterations :98-135)."""
import logging
import time

from .booster import LightGBMBooster

log = logging.getLogger("synapseml_b200.lightgbm")


def create_booster(train_params, train_dataset):
    """TrainUtils.createBooster: one native booster over the executor's dataset (modelString merge: later row)."""
    if train_params.generalParams.modelString:
        raise NotImplementedError("continued training from modelString (LGBM_BoosterMerge) is not built yet")
    return LightGBMBooster(train_dataset, train_params.to_string())


def update_one_iteration(train_params, booster, cur_iters):
    """TrainUtils.updateOneIteration: a native failure ends training quietly (TrainUtils.scala:89-95) but is logged loudly."""
    try:
        if train_params.objectiveParams.fobj is not None:
            raise NotImplementedError("custom objective (LGBM_BoosterUpdateOneIterCustom) is not built yet")
        return booster.update_one_iteration()
    except Exception as e:  # noqa: BLE001 - mirrors the reference's catch-all
        log.warning("LightGBM reached early termination on one task, stopping training on task. "
                    "This message should rarely occur. Inner exception: %s", e)
        return True


def execute_training_iterations(train_params, booster, measures=None):
    """TrainUtils.executeTrainingIterations: the boosting loop; returns the best iteration (None without validation data)."""
    delegate = train_params.delegate
    num_iterations = train_params.generalParams.numIterations
    learning_rate = train_params.generalParams.learningRate
    t0 = time.time()
    cur_iters = 0
    is_finished = False
    while not is_finished and cur_iters < num_iterations:
        if delegate is not None:
            delegate.before_train_iteration(cur_iters)
            new_lr = delegate.get_learning_rate(cur_iters, learning_rate)
            if new_lr != learning_rate:
                learning_rate = new_lr
                booster.reset_parameter("learning_rate=%s" % learning_rate)
        is_finished = update_one_iteration(train_params, booster, cur_iters)
        if delegate is not None:
            delegate.after_train_iteration(cur_iters, is_finished)
        cur_iters += 1
    if measures is not None:
        measures["trainingIterationsTime"] = time.time() - t0
        measures["iterations"] = cur_iters
    return None
