"""TEST INFRASTRUCTURE: minimal stand-in for the un-vendored third-party `diffusers` package
(requirements.txt:33 of the reference: "developed using version ==0.25.0"), so that the reference's
decoder modules import unmodified in the build container when golden vectors are generated.
Restates the 0.25-era semantics of the two classes the reference actually uses
(matcha/models/components/transformer.py:14-23,156,251-259,320-325); see SURVEY.md Appendix A.
Never imported by the product and never shipped to users."""
