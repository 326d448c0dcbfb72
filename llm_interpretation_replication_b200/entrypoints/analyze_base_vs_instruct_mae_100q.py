#!/usr/bin/env python
"""Entry point 5 -- mirrors reference survey_analysis/analyze_base_vs_instruct_mae_100q.py (the BASELINE
config-5 bootstrap: 10,000 resamples x n questions, seed 42 per call).

Host side (pandas, as in the reference): survey loading + exclusion criteria (:102-175), question text
extraction (:178-197), human averages (:200-209), model table with relative_prob = yes/(yes+no) and
fillna(0.5) (:212-222), prompt matching (:225-236), data validation (:239-263), per-model absolute errors in CSV
row order (:270-286).  Device side: bootstrap_mae_difference / bootstrap_overall_mae_difference (:289-414).
Output: base_vs_instruct_mae_results.csv with columns Model_Family, Base_MAE, Instruct_MAE, Difference, CI_Lower,
CI_Upper, P_Value, N_Questions (+ OVERALL row) (:656-688).

    python -m llm_interpretation_replication_b200.entrypoints.analyze_base_vs_instruct_mae_100q \
        --survey S1.csv S2.csv --models model_comparison_results.csv [--out DIR]
"""
from __future__ import annotations

import argparse
import os

import numpy as np
import pandas as pd

from .. import bootstrap

MODEL_FAMILIES = {   # reference :60-85
    "Falcon": {"base": "tiiuae/falcon-7b", "instruct": "tiiuae/falcon-7b-instruct"},
    "StableLM": {"base": "stabilityai/stablelm-base-alpha-7b", "instruct": "stabilityai/stablelm-tuned-alpha-7b"},
    "RedPajama": {"base": "togethercomputer/RedPajama-INCITE-7B-Base", "instruct": "togethercomputer/RedPajama-INCITE-7B-Instruct"},
    "BLOOM": {"base": "bigscience/bloom-7b1", "instruct": "bigscience/bloomz-7b1"},
    "Pythia-Dolly": {"base": "EleutherAI/pythia-6.9b", "instruct": "databricks/dolly-v2-7b"},
    "Mistral": {"base": "mistralai/Mistral-7B-v0.1", "instruct": "mistralai/Mistral-7B-Instruct-v0.2"},
}
MIN_STD_THRESHOLD = 0.01
MAX_NAN_FRACTION = 0.5


def load_and_clean_survey_data(filepaths):
    if isinstance(filepaths, str):
        filepaths = [filepaths]
    dfs = []
    for s, path in enumerate(filepaths, start=1):
        t = pd.read_csv(path)[2:].reset_index(drop=True)          # two Qualtrics description rows
        ren = {f"Q{g}_{q}": f"S{s}_Q{g}_{q}" for g in range(1, 6) for q in range(1, 12) if f"Q{g}_{q}" in t.columns}
        dfs.append(t.rename(columns=ren))
    df = pd.concat(dfs, ignore_index=True)
    df["Duration (in seconds)"] = pd.to_numeric(df["Duration (in seconds)"], errors="coerce")
    qcols = []
    for s in range(1, len(filepaths) + 1):
        for g in range(1, 6):
            for q in range(1, 12):
                c = f"S{s}_Q{g}_{q}"
                if c in df.columns:
                    qcols.append(c)
                    df[c] = pd.to_numeric(df[c], errors="coerce")
    return df, qcols


def apply_exclusion_criteria(df, qcols):
    """Same three filters as the reference (:143-175), vectorised: duration < 20% of the median; all substantive
    answers identical (needs more than one answer); any answered attention check (`_8`) different from 100."""
    df = df[df["Duration (in seconds)"] >= 0.2 * df["Duration (in seconds)"].median()]
    sub = df[[c for c in qcols if not c.endswith("_8")]]
    identical = (sub.notna().sum(axis=1) > 1) & (sub.nunique(axis=1, dropna=True) == 1)
    df = df[~identical]
    att = df[[c for c in qcols if c.endswith("_8")]]
    failed = (att.notna() & (att != 100)).any(axis=1)
    return df[~failed]


def extract_question_text(filepaths):
    mapping = {}
    for s, path in enumerate(filepaths, start=1):
        raw = pd.read_csv(path)
        headers = raw.iloc[0]
        for col in raw.columns:
            if col.startswith("Q") and "_" in col:
                text = headers[col]
                if pd.notna(text) and isinstance(text, str) and " - " in text:
                    mapping[f"S{s}_{col}"] = text.split(" - ")[-1].strip()
    return mapping


def calculate_human_averages(df, qcols):
    return {c: df[c].dropna().mean() / 100.0 for c in qcols if not c.endswith("_8") and df[c].notna().any()}


def load_model_data(path):
    df = pd.read_csv(path)
    df["relative_prob"] = (df["yes_prob"] / (df["yes_prob"] + df["no_prob"])).fillna(0.5)
    return df


def match_prompts_to_questions(question_mapping, model_df):
    text_to_id = {v: k for k, v in question_mapping.items()}
    return {p: text_to_id[p] for p in model_df["prompt"].unique() if p in text_to_id}


def validate_model_data(model_df, model_name):
    data = model_df[model_df["model"] == model_name]["relative_prob"]
    if len(data) == 0:
        return False, "No data found"
    nan_fraction = data.isna().sum() / len(data)
    if nan_fraction > MAX_NAN_FRACTION:
        return False, f"{nan_fraction * 100:.0f}% NaN values"
    valid = data.dropna()
    if len(valid) > 1 and valid.std() < MIN_STD_THRESHOLD:
        return False, f"Constant values (std={valid.std():.4f})"
    return True, "OK"


def calculate_mae_per_model(model_df, human_avgs, matches, model_name):
    errors = []
    for _, row in model_df[model_df["model"] == model_name].iterrows():   # CSV row order, as the reference
        q = matches.get(row["prompt"])
        if q is not None and q in human_avgs:
            errors.append(abs(row["relative_prob"] - human_avgs[q]))
    return (np.mean(errors), errors) if errors else (None, [])


def family_errors(survey_paths, models_csv):
    """Everything up to the bootstrap: {family: {'base': [...], 'instruct': [...]}} in MODEL_FAMILIES order."""
    survey_df, qcols = load_and_clean_survey_data(survey_paths)
    survey_df = apply_exclusion_criteria(survey_df, qcols)
    qmap = {k: v for k, v in extract_question_text(survey_paths).items() if not k.endswith("_8")}
    human = calculate_human_averages(survey_df, qcols)
    model_df = load_model_data(models_csv)
    matches = match_prompts_to_questions(qmap, model_df)
    out = {}
    for fam, m in MODEL_FAMILIES.items():
        if m["base"] not in model_df["model"].values or m["instruct"] not in model_df["model"].values:
            continue
        if not validate_model_data(model_df, m["base"])[0] or not validate_model_data(model_df, m["instruct"])[0]:
            continue
        bm, be = calculate_mae_per_model(model_df, human, matches, m["base"])
        im, ie = calculate_mae_per_model(model_df, human, matches, m["instruct"])
        if bm is None or im is None:
            continue
        out[fam] = {"base": be, "instruct": ie}
    return out, len(survey_df)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--survey", nargs="+", required=True)
    ap.add_argument("--models", required=True)
    ap.add_argument("--out", default=".")
    args = ap.parse_args(argv)
    errors, n_resp = family_errors(args.survey, args.models)
    rows = []
    for fam, e in errors.items():
        r = bootstrap.bootstrap_mae_difference(e["base"], e["instruct"])
        rows.append({"Model_Family": fam, "Base_MAE": r["base_mae"], "Instruct_MAE": r["instruct_mae"],
                     "Difference": r["observed_diff"], "CI_Lower": r["ci_lower"], "CI_Upper": r["ci_upper"],
                     "P_Value": r["p_value"], "N_Questions": r["n_questions"]})
    if errors:
        o = bootstrap.bootstrap_overall_mae_difference(errors)
        rows.append({"Model_Family": "OVERALL", "Base_MAE": o["base_mae"], "Instruct_MAE": o["instruct_mae"],
                     "Difference": o["observed_diff"], "CI_Lower": o["ci_lower"], "CI_Upper": o["ci_upper"],
                     "P_Value": o["p_value"], "N_Questions": o["n_questions"]})
    os.makedirs(args.out, exist_ok=True)
    df = pd.DataFrame(rows)
    df.to_csv(os.path.join(args.out, "base_vs_instruct_mae_results.csv"), index=False)
    print(f"respondents after exclusion: {n_resp}")
    print(df.to_string(index=False))
    return df


if __name__ == "__main__":
    main()
