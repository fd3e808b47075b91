"""TrainUtils: mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/TrainUtils.scala
(createBooster :16-29, updateOneIteration :77-96, executeTrainingIterations :98-135)."""
import logging
import time

from .booster import LightGBMBooster

log = logging.getLogger("synapseml_b200.lightgbm")


def create_booster(train_params, train_dataset):
    """TrainUtils.createBooster: one native booster over the executor's dataset (+ the trees of modelString in front, when given)."""
    booster = LightGBMBooster(train_dataset, train_params.to_string())
    if train_params.generalParams.modelString:       # TrainUtils.scala:22-24
        booster.merge_booster(train_params.generalParams.modelString)
    return booster


def update_one_iteration(train_params, booster, cur_iters):
    """TrainUtils.updateOneIteration: a native failure ends training quietly (TrainUtils.scala:89-95) but is logged loudly."""
    try:
        fobj = train_params.objectiveParams.fobj
        if fobj is not None:       # TrainUtils.scala:80-86: raw training scores -> user gradients -> UpdateOneIterCustom
            # innerPredict (LightGBMBooster.scala:316-330): [rows][classes] from the class-major native scores
            k = booster.num_classes()
            gradient, hessian = fobj.get_gradient(booster.inner_predict(0).reshape(k, -1).T, booster.train_dataset)
            return booster.update_one_iteration_custom(gradient, hessian)
        return booster.update_one_iteration()
    except Exception as e:  # noqa: BLE001 - mirrors the reference's catch-all
        if train_params.generalParams.numMachines > 1:
            # with several machines a task that stops alone would leave the others waiting for its histograms: the native side
            # has told them to give up (LGBM_BoosterUpdateOneIter), and the failure is raised here instead of being swallowed
            raise
        log.warning("LightGBM reached early termination on one task, stopping training on task. "
                    "This message should rarely occur. Inner exception: %s", e)
        return True


_LARGER_IS_BETTER = ("auc", "ndcg@", "map@", "average_precision")


class TrainingState:
    """PartitionTaskTrainingState: what the early-stopping bookkeeping keeps between iterations."""

    def __init__(self, eval_names):
        self.eval_names = eval_names
        self.iteration = 0
        self.is_finished = False
        self.best_iteration_result = None
        n = len(eval_names)
        self.best_score = [None] * n
        self.best_iteration = [0] * n
        self.best_scores = [None] * n


def get_valid_eval_results(state, booster, early_stopping_round, improvement_tolerance):
    """TrainUtils.getValidEvalResults (:143-169): per metric, remember the best value; stop when a metric has not improved by more
    than the tolerance for earlyStoppingRound iterations.  (As in the reference the check runs for earlyStoppingRound = 0 too.)"""
    results = booster.get_eval_results(state.eval_names, 1)
    for index, (name, score) in enumerate(results):
        log.info("Valid %s=%s", name, score)
        if name.startswith(_LARGER_IS_BETTER):
            improved = state.best_scores[index] is None or score - state.best_score[index] > improvement_tolerance
        else:
            improved = state.best_scores[index] is None or score - state.best_score[index] < improvement_tolerance
        if improved:
            state.best_score[index] = score
            state.best_iteration[index] = state.iteration
            state.best_scores[index] = [v for _, v in results]
        elif state.iteration - state.best_iteration[index] >= early_stopping_round:
            state.is_finished = True
            log.info("Early stopping, best iteration is %d", state.best_iteration[index])
            state.best_iteration_result = state.best_iteration[index]
    return dict(results)


def execute_training_iterations(train_params, booster, measures=None, has_validation_data=False):
    """TrainUtils.executeTrainingIterations (:98-135): the boosting loop; returns the best iteration (None unless early stopping
    triggered)."""
    delegate = train_params.delegate
    gp = train_params.generalParams
    num_iterations = gp.numIterations
    learning_rate = gp.learningRate
    provide_train_metric = bool(train_params.isProvideTrainingMetric)
    state = TrainingState(booster.get_eval_names() if (has_validation_data or provide_train_metric) else [])
    history = []
    t0 = time.time()
    while True:
        if delegate is not None:
            delegate.before_train_iteration(state.iteration)
            new_lr = delegate.get_learning_rate(state.iteration, learning_rate)
            if new_lr != learning_rate:
                learning_rate = new_lr
                booster.reset_parameter("learning_rate=%s" % learning_rate)
        state.is_finished = update_one_iteration(train_params, booster, state.iteration)
        train_eval = valid_eval = None
        if provide_train_metric and not state.is_finished:
            train_eval = dict(booster.get_eval_results(state.eval_names, 0))
            for k, v in train_eval.items():
                log.info("Train %s=%s", k, v)
        if has_validation_data and not state.is_finished:
            valid_eval = get_valid_eval_results(state, booster, gp.earlyStoppingRound, gp.improvementTolerance)
        history.append((train_eval, valid_eval))
        if delegate is not None:
            delegate.after_train_iteration(state.iteration, state.is_finished)
        state.iteration += 1
        if state.is_finished or state.iteration >= num_iterations:
            break
    if measures is not None:
        measures["trainingIterationsTime"] = time.time() - t0
        measures["iterations"] = state.iteration
        measures["evalHistory"] = history
    return state.best_iteration_result
