#!/usr/bin/env python
"""Synthesise the two JSON inputs reference survey_analysis/bootstrap_confidence_intervals.py reads (:9-20) but the
reference never produces in a usable form: `consolidated_analysis_results.json`
(`matching_stats.matches`: prompt -> "Q{g}_{k}") and `survey_analysis_detailed.json`
(`results.by_question[qid]`: mean_response / std_response on the 0-100 scale).  Both are derived from the shipped
survey-1 CSV exactly as the reference derives its human statistics (cleaning + exclusion of
analyze_base_vs_instruct_mae_100q.py:102-175; np.mean / np.std of survey_analysis_consolidated.py:179-194).
Writes tests/golden/ci_fixture.json.  Needs only the CSV fixtures next to this file."""
import json
import os
import sys

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from llm_interpretation_replication_b200.entrypoints import analyze_base_vs_instruct_mae_100q as E5  # noqa: E402

survey = [os.path.join(HERE, "word_meaning_survey_results.csv")]
df, qcols = E5.load_and_clean_survey_data(survey)
df = E5.apply_exclusion_criteria(df, qcols)
qmap = {k: v for k, v in E5.extract_question_text(survey).items() if not k.endswith("_8")}
model_df = pd.read_csv(os.path.join(HERE, "model_comparison_results.csv"))
matches = E5.match_prompts_to_questions(qmap, model_df)                     # prompt -> "S1_Qg_k"
by_q = {}
for col in qcols:
    if col.endswith("_8"):
        continue
    r = df[col].dropna()
    if len(r):
        by_q[col.replace("S1_", "")] = {"mean_response": float(np.mean(r)), "std_response": float(np.std(r)), "n": int(len(r))}
out = {"consolidated": {"matching_stats": {"matches": {p: q.replace("S1_", "") for p, q in matches.items()}}},
       "detailed": {"results": {"by_question": by_q}}}
json.dump(out, open(os.path.join(HERE, "ci_fixture.json"), "w"), indent=1)
print(len(matches), "matched prompts;", len(by_q), "questions with statistics")
