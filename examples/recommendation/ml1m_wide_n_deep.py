"""Wide&Deep on MovieLens-1M-shaped data with the recommender entry points (reference example:
zoo/.../examples/recommendation/Ml1mWideAndDeep.scala: ratings / users / movies joined, gender + occupation as wide base columns,
age x gender as a crossed column, genres + gender as indicators, userId + itemId as embeddings, age continuous, 5 rating classes;
`fit`, `predict_user_item_pair`, `recommend_for_user`, `recommend_for_item`).

The three tables are generated (no network) with ML-1M's schema: a user's rating depends on whether the movie's genre is one of the
genres preferred by the user's (gender, age) group, so the crossed and indicator columns carry the signal."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd

from zoo.common.nncontext import init_nncontext
from zoo.models.recommendation import ColumnFeatureInfo, WideAndDeep
from zoo.models.recommendation.utils import categorical_from_vocab_list, hash_bucket, to_user_item_feature
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.api.keras.optimizers import Adam

GENRES = ["Crime", "Romance", "Thriller", "Adventure", "Drama", "Children's", "War", "Documentary", "Fantasy", "Mystery", "Musical",
          "Animation", "Film-Noir", "Horror", "Western", "Comedy", "Action", "Sci-Fi"]


def load_public_data(users, items, ratings, seed=0):
    """Tables with the schema of users.dat / movies.dat / ratings.dat."""
    rng = np.random.RandomState(seed)
    user_df = pd.DataFrame({"userId": np.arange(1, users + 1), "gender": rng.choice(["F", "M"], users),
                            "age": rng.choice([1, 18, 25, 35, 45, 50, 56], users), "occupation": rng.randint(0, 21, users)})
    item_df = pd.DataFrame({"itemId": np.arange(1, items + 1), "title": ["movie %d" % i for i in range(1, items + 1)],
                            "genres": rng.choice(GENRES, items)})
    r = pd.DataFrame({"userId": rng.randint(1, users + 1, ratings), "itemId": rng.randint(1, items + 1, ratings)}).drop_duplicates()
    j = r.merge(user_df, on="userId").merge(item_df, on="itemId")
    group = (j["gender"] == "M").astype(int).values * 7 + pd.Categorical(j["age"]).codes
    liked = (np.array([GENRES.index(g) for g in j["genres"]]) + group) % 3 == 0          # each (gender, age) group likes a third of the genres
    j["label"] = np.clip(np.where(liked, 4, 2) + rng.randint(0, 2, len(j)), 1, 5)
    return j[["userId", "itemId", "label"]], user_df, item_df


def assembly_feature(ratings, user_df, item_df, column_info, model_type):
    users = user_df.assign(**{"age-gender": [hash_bucket("%s_%s" % (a, g), 100) for a, g in zip(user_df["age"], user_df["gender"])],
                              "gender": [categorical_from_vocab_list(g, ["F", "M"], start=1) for g in user_df["gender"]]})
    items = item_df.assign(genres=[categorical_from_vocab_list(g, GENRES, start=1) for g in item_df["genres"]])
    joined = ratings.merge(items, on="itemId").merge(users, on="userId")
    return [to_user_item_feature(row, column_info, model_type) for row in joined.to_dict("records")]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model-type", default="wide_n_deep", choices=["wide", "deep", "wide_n_deep"])
    ap.add_argument("--epochs", type=int, default=int(os.environ.get("EPOCHS", 6)))
    ap.add_argument("--batch-size", type=int, default=512)
    ap.add_argument("--ratings", type=int, default=12000)
    a = ap.parse_args()
    init_nncontext("WideAndDeepExample")
    ratings, user_df, item_df = load_public_data(300, 200, a.ratings)
    print(ratings.groupby("label").size().to_string())
    info = ColumnFeatureInfo(wide_base_cols=["occupation", "gender"], wide_base_dims=[21, 3], wide_cross_cols=["age-gender"], wide_cross_dims=[100],
                             indicator_cols=["genres", "gender"], indicator_dims=[19, 3], embed_cols=["userId", "itemId"],
                             embed_in_dims=[301, 201], embed_out_dims=[64, 64], continuous_cols=["age"])
    wnd = WideAndDeep(5, info, model_type=a.model_type)
    feats = assembly_feature(ratings, user_df, item_df, info, a.model_type)
    rng = np.random.RandomState(1)
    order = rng.permutation(len(feats))
    cut = int(0.8 * len(feats))
    train, val = [feats[i] for i in order[:cut]], [feats[i] for i in order[cut:]]
    wnd.compile(optimizer=Adam(lr=1e-2, decay=1e-5), loss=SparseCategoricalCrossEntropy(zero_based_label=False), metrics=["accuracy"])
    wnd.fit([f.sample for f in train], batch_size=a.batch_size, nb_epoch=a.epochs, validation_data=[f.sample for f in val])
    print("validation:", wnd.evaluate([f.sample for f in val], batch_size=a.batch_size))
    for p in wnd.predict_user_item_pair(val)[:5]:
        print(p)
    for rec in wnd.recommend_for_user(val, 3)[:3] + wnd.recommend_for_item(val, 3)[:3]:
        print(rec)
    print("finished...")
