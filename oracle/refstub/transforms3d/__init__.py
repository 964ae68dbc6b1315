"""TEST INFRASTRUCTURE -- placeholder so that the reference's package __init__ imports resolve
(BetaAviary / CTBRControl are out of scope and never executed)."""
