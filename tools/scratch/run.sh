bash tools/final_evidence.sh v52
