"""Stand-in for the ``xgboost`` package, used ONLY by ``bench.py --impl reference``.

The reference's ``scripts/model_comp.py`` does ``from xgboost import XGBClassifier`` at import time; xgboost is not
installed in this image and there is no wheel in the offline wheelhouse, so the module could not even be imported.
This package satisfies that one import with scikit-learn's ``GradientBoostingClassifier`` carrying the
hyper-parameters the reference passes (learning_rate, max_depth, n_estimators, random_state).

scikit-learn's booster is far slower than xgboost's C++ core on this shape (about 30 s for the 700 trees against what
would be about a second), so the stand-in TIMES ITSELF: ``seconds_inside`` accumulates the wall-clock spent in its
``fit`` / ``predict_proba``, and the reference arm reports its time with that share taken out - the reference gets its
xgboost stage for free rather than being charged for a substitute.  It is a dependency substitute for the reference
arm, declared in the arm's JSON line; nothing of consensusml_b200 is involved and the framework never imports it."""
import time

from sklearn.ensemble import GradientBoostingClassifier

seconds_inside = 0.0


class XGBClassifier(GradientBoostingClassifier):
    def __init__(self, learning_rate=0.3, max_depth=6, n_estimators=100, random_state=None, **ignored):
        super().__init__(learning_rate=learning_rate, max_depth=max_depth, n_estimators=n_estimators, random_state=random_state)

    def fit(self, X, y, **kw):
        global seconds_inside
        t0 = time.perf_counter()
        try:
            return super().fit(X, y, **kw)
        finally:
            seconds_inside += time.perf_counter() - t0

    def predict_proba(self, X):
        global seconds_inside
        t0 = time.perf_counter()
        try:
            return super().predict_proba(X)
        finally:
            seconds_inside += time.perf_counter() - t0


__version__ = "0-standin-sklearn"
