bash tools/final_evidence.sh v46
